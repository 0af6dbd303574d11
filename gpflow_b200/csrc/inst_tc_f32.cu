// explicit instantiation of the chunked tensor-core path (gpflow_tc*.cuh): fp32, hidden N tile 64 or 256
#include "gpflow_tc_ops.cuh"

namespace gpflow {

bool tc_match(const GpFlowDesc* d) { return tc::tc_shape_ok(d); }
size_t tc_workspace(const GpFlowDesc* d, int64_t B, int op) { return tc::tc_workspace_bytes(d, B, op); }
unsigned long long tc_launches_total() { return tc::tc_launch_counter(); }
int tc_slab(const GpFlowDesc* d, int64_t B, int op) { return tc::tc_slab_size(d, B, op); }

int tc_f32(const GpFlowDesc* d, int op, const float* params, const float* x_in, const float* x_data, const float* grad_y,
           float* y_out, float* logdet_out, float* traj_out, float* grad_x_out, float* grads_out, float* scalars_out, int64_t B,
           void* ws, size_t wsb, cudaStream_t st, int sm_count, float* v_out, double t_eval) {
  if (!tc::tc_shape_ok(d)) return GPFLOW_ERR_UNSUPPORTED;
  const tc::TcDims g = tc::tc_make_dims(d, 256);
  if (g.HT == 64)
    return tc::tc_run<64, 4>(d, op, params, x_in, x_data, grad_y, y_out, logdet_out, traj_out, grad_x_out, grads_out, scalars_out, B,
                             ws, wsb, st, sm_count, v_out, t_eval);
  return tc::tc_run<256, 2>(d, op, params, x_in, x_data, grad_y, y_out, logdet_out, traj_out, grad_x_out, grads_out, scalars_out, B,
                            ws, wsb, st, sm_count, v_out, t_eval);
}

}  // namespace gpflow
