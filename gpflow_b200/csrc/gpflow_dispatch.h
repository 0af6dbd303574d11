// gpflow_b200 -- internal: entry points of the per-(dtype, euler) translation units (compiled in parallel).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/gpflow_b200.h"

namespace gpflow {

constexpr int kSmallBlock = 128;
constexpr int kMaxSMs = 160;  // upper bound used for device-independent workspace sizing (B200: 148)

template <typename T>
struct Ptrs {
  const T* params; const T* x_in; const T* x_data; const T* grad_y; const T* reg_cot; const T* zw;
  T* y_out; T* logdet_out; T* traj_out; T* grad_x_out; T* grads_out; T* scalars_out;
};

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// small path workspace: [partials grid_max*npart][traj (N+1)*d*TS_max]
inline size_t small_workspace(const GpFlowDesc* d, int64_t B, int op, int nparam) {
  const size_t es = d->dtype == GPFLOW_F64 ? 8 : 4;
  const int64_t max_threads = (int64_t)kMaxSMs * 2048;
  int64_t ts = (int64_t)align_up((size_t)B, kSmallBlock);
  if (ts > max_threads) ts = max_threads;
  const int64_t grid_max = ts / kSmallBlock;
  size_t bytes = align_up((size_t)grid_max * (3 + nparam) * es, 256);
  if (op != GPFLOW_OP_FORWARD) bytes += align_up((size_t)(d->n_steps + 1) * d->d * ts * es, 256);
  return bytes;
}

// one function per (shape, dtype, euler) translation unit; returns GPFLOW_ERR_UNSUPPORTED if not compiled in
#define GPF_DECL_SMALL(NAME) \
  int NAME(const GpFlowDesc* d, int op, const void* ptrs, int64_t B, void* ws, size_t wsb, cudaStream_t st, int sm_count)
GPF_DECL_SMALL(small_2_1_15_15_f32);
GPF_DECL_SMALL(small_2_1_15_15_f32_euler);
GPF_DECL_SMALL(small_2_1_15_15_f64);
GPF_DECL_SMALL(small_2_1_15_15_f64_euler);
GPF_DECL_SMALL(sf2_15_15_f32);  // stream-function kernels v2 (fp32, no Euler): train step / backward, forward-only
GPF_DECL_SMALL(sf2e_15_15_f32);  // sf2r + Euler penalization (train step / backward)
bool sf2e_15_15_enabled();
size_t sf2e_15_15_workspace(const GpFlowDesc* d, int64_t B);
size_t sf2_15_15_workspace(const GpFlowDesc* d, int64_t B, int op);
int sf2_15_15_launches(int op);              // [pack +] main kernel + finalize
const char* sf2_15_15_kernel_name(int op);   // gpflow_sf2r_kernel (train / backward, default) or gpflow_sf2_kernel

#define GPF_DECL_GENERIC(NAME, T)                                                                                       \
  int NAME(const GpFlowDesc* d, int op, const T* params, const T* x_in, const T* x_data, const T* grad_y,               \
           const T* reg_cot, const T* zw, T* y_out, T* logdet_out, T* traj_out, T* grad_x_out, T* grads_out,            \
           T* scalars_out, int64_t B, void* ws, size_t wsb, cudaStream_t st, int sm_count, T* v_out, double t_eval)
GPF_DECL_GENERIC(generic_f32, float);
GPF_DECL_GENERIC(generic_f32_euler, float);
GPF_DECL_GENERIC(generic_f64, double);
GPF_DECL_GENERIC(generic_f64_euler, double);

// chunked tensor-core path (gpflow_tc*.cuh): fp32, no Euler penalization, d <= 64, hidden widths <= 256
bool tc_match(const GpFlowDesc* d);
size_t tc_workspace(const GpFlowDesc* d, int64_t B, int op);
unsigned long long tc_launches_total();
int tc_slab(const GpFlowDesc* d, int64_t B, int op);
int tc_f32(const GpFlowDesc* d, int op, const float* params, const float* x_in, const float* x_data, const float* grad_y,
           float* y_out, float* logdet_out, float* traj_out, float* grad_x_out, float* grads_out, float* scalars_out, int64_t B,
           void* ws, size_t wsb, cudaStream_t st, int sm_count, float* v_out, double t_eval);

size_t generic_workspace_bytes(const GpFlowDesc* d, int64_t B, int op);
bool generic_runs_padded(const GpFlowDesc* d);   // hidden widths zero-padded to the aligned mma shape (one more launch)

}  // namespace gpflow
