// translation unit: sf2 (2-D stream-function) kernels, float: train step / backward (TMA-stashed) and forward-only
#include "gpflow_sf2_fwd.cuh"
namespace gpflow {
GPF_DECL_SMALL(sf2_15_15_f32) {
  const Ptrs<float>& p = *reinterpret_cast<const Ptrs<float>*>(ptrs);
  if (op == GPFLOW_OP_FORWARD) return launch_sf2_fwd<15, 15>(d, p, B, ws, wsb, st, sm_count);
  return launch_sf2<15, 15>(d, op, p, B, ws, wsb, st, sm_count);
}
size_t sf2_15_15_workspace(const GpFlowDesc* d, int64_t B, int op) {
  return op == GPFLOW_OP_FORWARD ? sf2_fwd_workspace<15, 15>(B) : sf2_workspace<15, 15>(d, B);
}
}  // namespace gpflow
