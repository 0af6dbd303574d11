// translation unit: sf2 (2-D stream-function) train-step kernel, float
#include "gpflow_sf2_kernel.cuh"
namespace gpflow {
GPF_DECL_SMALL(sf2_15_15_f32) {
  (void)wsb;
  return launch_sf2<15, 15>(d, op, *reinterpret_cast<const Ptrs<float>*>(ptrs), B, ws, wsb, st, sm_count);
}
size_t sf2_15_15_workspace(const GpFlowDesc* d, int64_t B) { return sf2_workspace<15, 15>(d, B); }
}  // namespace gpflow
