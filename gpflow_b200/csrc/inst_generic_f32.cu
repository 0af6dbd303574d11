// translation unit: generic kernels, float
#include "gpflow_dispatch.h"
#include "gpflow_generic.cuh"
namespace gpflow {
GPF_DECL_GENERIC(generic_f32, float) {
  return launch_generic_impl<float, false>(d, op, params, x_in, x_data, grad_y, reg_cot, zw, y_out, logdet_out, traj_out,
                                      grad_x_out, grads_out, scalars_out, B, ws, wsb, st, sm_count, v_out, t_eval);
}
size_t generic_workspace_bytes(const GpFlowDesc* d, int64_t B, int op) { return generic_workspace(d, B, op); }
bool generic_runs_padded(const GpFlowDesc* d) {
  GpFlowDesc dp;
  return generic_padded_desc(d, &dp);
}
}  // namespace gpflow
