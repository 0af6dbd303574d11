// translation unit: small-d kernels, float
#include "gpflow_small_launch.cuh"
namespace gpflow {
GPF_DECL_SMALL(small_2_1_15_15_f32) { return launch_small_shape<float, 2, 1, 15, 15, false>(d, op, ptrs, B, ws, wsb, st, sm_count); }
}  // namespace gpflow
