// translation unit: generic kernels, double, Euler penalization
#include "gpflow_dispatch.h"
#include "gpflow_generic.cuh"
namespace gpflow {
GPF_DECL_GENERIC(generic_f64_euler, double) {
  return launch_generic_impl<double, true>(d, op, params, x_in, x_data, grad_y, reg_cot, zw, y_out, logdet_out, traj_out,
                                      grad_x_out, grads_out, scalars_out, B, ws, wsb, st, sm_count, v_out, t_eval);
}
}  // namespace gpflow
