// translation unit: small-d kernels, double, Euler penalization
#include "gpflow_small_launch.cuh"
namespace gpflow {
GPF_DECL_SMALL(small_2_1_15_15_f64_euler) { return launch_small_shape<double, 2, 1, 15, 15, true>(d, op, ptrs, B, ws, wsb, st, sm_count); }
}  // namespace gpflow
