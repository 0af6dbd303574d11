// gpflow_b200 -- C ABI (include/gpflow_b200.h): descriptor validation, kernel dispatch, workspace carving.
// No torch types, no allocation, no synchronisation: everything is enqueued on the caller's stream.
#include "../../include/gpflow_b200.h"

#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>
#include <math.h>
#include <stdlib.h>

#include <mutex>

#include "gpflow_dispatch.h"
#include "gpflow_bnaf.cuh"
#include "gpflow_ffjord.cuh"

using namespace gpflow;

namespace {

struct DeviceInfo {
  int sm_count = 148;
  bool ok = false;
};
DeviceInfo device_info() {
  static std::mutex mu;
  static DeviceInfo cache[64];
  static bool have[64] = {false};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
    cudaGetLastError();
    return DeviceInfo();
  }
  std::lock_guard<std::mutex> lk(mu);
  if (!have[dev]) {
    DeviceInfo di;
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0) {
      di.sm_count = v > kMaxSMs ? kMaxSMs : v;
      di.ok = true;
    }
    cudaGetLastError();
    cache[dev] = di;
    have[dev] = true;
  }
  return cache[dev];
}

int validate(const GpFlowDesc* d) {
  if (d == nullptr) return GPFLOW_ERR_NULL_POINTER;
  if (d->abi_version != GPFLOW_ABI_VERSION) return GPFLOW_ERR_BAD_DESC;
  if (d->dtype != GPFLOW_F32 && d->dtype != GPFLOW_F64) return GPFLOW_ERR_BAD_DESC;
  if (d->d < 1 || d->nb < 1 || d->nb > d->d) return GPFLOW_ERR_BAD_DESC;
  if (d->n_hidden < 1 || d->n_hidden > GPFLOW_MAX_HIDDEN) return GPFLOW_ERR_BAD_DESC;
  for (int i = 0; i < d->n_hidden; ++i)
    if (d->hidden[i] < 1) return GPFLOW_ERR_BAD_DESC;
  if (d->n_steps < 1 || d->n_steps_nominal < 1) return GPFLOW_ERR_BAD_DESC;
  if (!(d->dt > 0.0) || !(d->sigma > 0.0)) return GPFLOW_ERR_BAD_DESC;
  if (d->d < 2) return GPFLOW_ERR_UNSUPPORTED;  // a 1-d divergence-free wall-tangent field is identically zero
  return GPFLOW_OK;
}

// ---- the specialised small-d variants that are compiled in -------------------------------------
struct SmallShape { int d, nb, h1, h2; };
bool small_match(const GpFlowDesc* d, SmallShape* out) {
  if (d->flags & GPFLOW_FLAG_FORCE_GENERIC) return false;
  if (d->n_hidden != 2) return false;
  static const SmallShape shapes[] = {{2, 1, 15, 15}};
  for (const SmallShape& s : shapes)
    if (d->d == s.d && d->nb == s.nb && d->hidden[0] == s.h1 && d->hidden[1] == s.h2) {
      if (out) *out = s;
      return true;
    }
  return false;
}

// sf2 (stream-function kernels v2) covers fp32 ops of the (2,1,15,15) shape without Euler penalization: forward
// (with or without the Lagrangian diagnostic / trajectory), train step and backward (without the Lagrangian);
// everything else of that shape stays on the v1 kernel.
bool sf2_match(const GpFlowDesc* d, int op) {
  SmallShape s;
  if (!small_match(d, &s) || d->dtype != GPFLOW_F32) return false;
  if (d->flags & (GPFLOW_FLAG_EULER_PEN | GPFLOW_FLAG_FORCE_V1)) return false;
  if (op != GPFLOW_OP_FORWARD && (d->flags & GPFLOW_FLAG_LAGRANGIAN)) return false;  // diagnostic: forward kernel only
  return d->n_steps <= 1024;
}

// sf2e: the same shape WITH the Euler penalization, train step / backward op (forward op, Lagrangian: v1 kernel)
bool sf2e_match(const GpFlowDesc* d, int op) {
  SmallShape s;
  if (!small_match(d, &s) || d->dtype != GPFLOW_F32 || !sf2e_15_15_enabled()) return false;
  if (!(d->flags & GPFLOW_FLAG_EULER_PEN) || (d->flags & (GPFLOW_FLAG_FORCE_V1 | GPFLOW_FLAG_LAGRANGIAN))) return false;
  return op != GPFLOW_OP_FORWARD && d->n_steps <= 512;   // 8 x 28 KB of warp buffers + the time table must fit 227 KB
}

// Narrower 2-D nets (d = 2, nb = 1, two hidden layers of width <= 15, fp32) run on the same [15,15] kernels with the
// parameters ZERO-PADDED to that shape: a padded unit has a = 0, h = tanh 0 = 0, and every weight that leaves it is 0, so
// it contributes nothing to psi, grad psi or any kept gradient entry.  Two tiny kernels around the call re-lay the flat
// parameter vector ([H1][3] | [H1] | [H2][H1] | [H2] | [2][H2] | [2], vector_field.py:44-73) and gather the gradient back.
// Without this, e.g. H = [10,10] fell to the generic kernel (30 x slower).  Shapes the sf2 / sf2e kernels do not take at
// [15,15] (fp64, Lagrangian in the train step, FORCE_V1) stay on the generic kernel.
bool pad2d_match(const GpFlowDesc* d, int op, GpFlowDesc* d15) {
  if (d->flags & (GPFLOW_FLAG_FORCE_GENERIC | GPFLOW_FLAG_FORCE_TC)) return false;
  if (d->dtype != GPFLOW_F32 || d->d != 2 || d->nb != 1 || d->n_hidden != 2) return false;
  if (d->hidden[0] > 15 || d->hidden[1] > 15 || (d->hidden[0] == 15 && d->hidden[1] == 15)) return false;
  GpFlowDesc t = *d;
  t.hidden[0] = t.hidden[1] = 15;
  if (!sf2_match(&t, op) && !sf2e_match(&t, op)) return false;
  if (d15) *d15 = t;
  return true;
}
constexpr int kPad2dParams = 332;                 // parameter count of the [15,15] shape
constexpr size_t kPad2dHead = 2 * 1536;           // padded parameters | padded gradient, 256-byte aligned
// index of entry e of the (h1, h2) layout in the [15,15] layout
__host__ __device__ inline int pad2d_index(int e, int h1, int h2) {
  const int o_b1 = 3 * h1, o_w2 = 4 * h1, o_b2 = o_w2 + h2 * h1, o_w3 = o_b2 + h2, o_b3 = o_w3 + 2 * h2;
  if (e < o_b1) return e;
  if (e < o_w2) return 45 + (e - o_b1);
  if (e < o_b2) return 60 + ((e - o_w2) / h1) * 15 + (e - o_w2) % h1;
  if (e < o_w3) return 285 + (e - o_b2);
  if (e < o_b3) return 300 + ((e - o_w3) / h2) * 15 + (e - o_w3) % h2;
  return 330 + (e - o_b3);
}
__global__ void pad2d_scatter_kernel(const float* __restrict__ src, float* __restrict__ dst15, int h1, int h2, int n) {
  __shared__ float buf[kPad2dParams];
  for (int i = threadIdx.x; i < kPad2dParams; i += blockDim.x) buf[i] = 0.0f;
  __syncthreads();
  for (int e = threadIdx.x; e < n; e += blockDim.x) buf[pad2d_index(e, h1, h2)] = src[e];
  __syncthreads();
  for (int i = threadIdx.x; i < kPad2dParams; i += blockDim.x) dst15[i] = buf[i];
}
__global__ void pad2d_gather_kernel(const float* __restrict__ src15, float* __restrict__ dst, int h1, int h2, int n) {
  for (int e = threadIdx.x; e < n; e += blockDim.x) dst[e] = src15[pad2d_index(e, h1, h2)];
}
size_t pad2d_workspace(const GpFlowDesc* d15, int64_t B, int op) {
  return kPad2dHead + (sf2_match(d15, op) ? sf2_15_15_workspace(d15, B, op) : sf2e_15_15_workspace(d15, B));
}
int pad2d_run(const GpFlowDesc* d, const GpFlowDesc* d15, int op, const Ptrs<float>& p, int64_t B, void* ws, size_t wsb,
              cudaStream_t st, int sm) {
  if (ws == nullptr || wsb < pad2d_workspace(d15, B, op)) return GPFLOW_ERR_WORKSPACE;
  const int n = (int)gpflow_param_count(d);
  float* p15 = reinterpret_cast<float*>(ws);
  float* g15 = reinterpret_cast<float*>(reinterpret_cast<char*>(ws) + kPad2dHead / 2);
  pad2d_scatter_kernel<<<1, 256, 0, st>>>(p.params, p15, d->hidden[0], d->hidden[1], n);
  if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  Ptrs<float> q = p;
  q.params = p15;
  if (p.grads_out != nullptr) q.grads_out = g15;
  void* inner = reinterpret_cast<char*>(ws) + kPad2dHead;
  const int rc = sf2_match(d15, op) ? sf2_15_15_f32(d15, op, &q, B, inner, wsb - kPad2dHead, st, sm)
                                    : sf2e_15_15_f32(d15, op, &q, B, inner, wsb - kPad2dHead, st, sm);
  if (rc != GPFLOW_OK) return rc;
  if (p.grads_out != nullptr) {
    pad2d_gather_kernel<<<1, 256, 0, st>>>(g15, p.grads_out, d->hidden[0], d->hidden[1], n);
    if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  }
  return GPFLOW_OK;
}

// tc (chunked tensor-core path) takes the wide fp32 shapes without Euler penalization once the batch is large enough to
// fill 128-row MMA tiles on every SM; GPFLOW_TC_MIN_BATCH overrides the threshold, GPFLOW_FLAG_FORCE_TC ignores it.
// Measured crossovers (B200, train step, profiles/r2_tc_crossover.txt): d=10, H=[50,50,50] -- generic 29.6 ms vs tc 33.5 ms at
// B = 4096, 58.4 vs 47.2 ms at 8192; d=64, H=[256,256] -- tc 8x faster already at B = 256.  Re-measured after the narrow-width
// work on the tc path (tools/tc_crossover.sh, profiles/r2_tc_crossover_v2.txt): d=10 -- generic 14.9 / 28.5 / 29.6 / 44.1 / 58.5 ms
// against tc 19.3 / 20.6 / 23.5 / 27.7 / 32.1 ms at B = 2048 / 3072 / 4096 / 6144 / 8192 (the generic kernel steps up with every
// 2368 samples = one round of its warp slots).
int64_t tc_min_batch(const GpFlowDesc* d) {
  static const int64_t env = [] { const char* e = getenv("GPFLOW_TC_MIN_BATCH"); return e ? (int64_t)atoll(e) : (int64_t)-1; }();
  if (env >= 0) return env;
  int hmax = 0;
  for (int i = 0; i < d->n_hidden; ++i) hmax = d->hidden[i] > hmax ? d->hidden[i] : hmax;
  return hmax <= 64 ? 2400 : 128;   // narrow nets: from the batch that needs a second round of the generic kernel's 148 x 16 warp slots
}
bool tc_selected(const GpFlowDesc* d, int64_t B) {
  if (d->flags & (GPFLOW_FLAG_FORCE_GENERIC | GPFLOW_FLAG_EULER_PEN)) return false;
  if (small_match(d, nullptr) || !tc_match(d)) return false;
  return (d->flags & GPFLOW_FLAG_FORCE_TC) != 0 || B >= tc_min_batch(d);
}

template <typename T>
int dispatch(const GpFlowDesc* d, int op, const Ptrs<T>& p, int64_t B, void* ws, size_t wsb, cudaStream_t st);

template <>
int dispatch<float>(const GpFlowDesc* d, int op, const Ptrs<float>& p, int64_t B, void* ws, size_t wsb, cudaStream_t st) {
  const int sm = device_info().sm_count;
  const bool euler = (d->flags & GPFLOW_FLAG_EULER_PEN) != 0;
  SmallShape s;
  if (sf2_match(d, op)) return sf2_15_15_f32(d, op, &p, B, ws, wsb, st, sm);
  if (sf2e_match(d, op)) return sf2e_15_15_f32(d, op, &p, B, ws, wsb, st, sm);
  if (small_match(d, &s)) return euler ? small_2_1_15_15_f32_euler(d, op, &p, B, ws, wsb, st, sm) : small_2_1_15_15_f32(d, op, &p, B, ws, wsb, st, sm);
  GpFlowDesc d15;
  if (pad2d_match(d, op, &d15)) return pad2d_run(d, &d15, op, p, B, ws, wsb, st, sm);
  if (tc_selected(d, B))
    return tc_f32(d, op, p.params, p.x_in, p.x_data, p.grad_y, p.y_out, p.logdet_out, p.traj_out, p.grad_x_out, p.grads_out,
                  p.scalars_out, B, ws, wsb, st, sm, nullptr, 0.0);
  if ((d->flags & GPFLOW_FLAG_FORCE_TC) && !(d->flags & GPFLOW_FLAG_FORCE_GENERIC)) return GPFLOW_ERR_UNSUPPORTED;
  return (euler ? generic_f32_euler : generic_f32)(d, op, p.params, p.x_in, p.x_data, p.grad_y, p.reg_cot, p.zw, p.y_out,
                                                 p.logdet_out, p.traj_out, p.grad_x_out, p.grads_out, p.scalars_out, B, ws,
                                                 wsb, st, sm, nullptr, 0.0);
}
template <>
int dispatch<double>(const GpFlowDesc* d, int op, const Ptrs<double>& p, int64_t B, void* ws, size_t wsb, cudaStream_t st) {
  const int sm = device_info().sm_count;
  const bool euler = (d->flags & GPFLOW_FLAG_EULER_PEN) != 0;
  SmallShape s;
  if (small_match(d, &s)) return euler ? small_2_1_15_15_f64_euler(d, op, &p, B, ws, wsb, st, sm) : small_2_1_15_15_f64(d, op, &p, B, ws, wsb, st, sm);
  return (euler ? generic_f64_euler : generic_f64)(d, op, p.params, p.x_in, p.x_data, p.grad_y, p.reg_cot, p.zw, p.y_out,
                                                 p.logdet_out, p.traj_out, p.grad_x_out, p.grads_out, p.scalars_out, B, ws,
                                                 wsb, st, sm, nullptr, 0.0);
}

bool misaligned(const void* p, size_t a) { return p != nullptr && (reinterpret_cast<uintptr_t>(p) % a) != 0; }

int run(const GpFlowDesc* d, int op, const void* params, const void* x_in, const void* x_data, const void* grad_y,
        const void* reg_cot, const void* zw, void* y_out, void* logdet_out, void* traj_out, void* grad_x_out,
        void* grads_out, void* scalars_out, int64_t B, void* ws, size_t wsb, void* stream) {
  const int rc = validate(d);
  if (rc != GPFLOW_OK) return rc;
  if (B < 0) return GPFLOW_ERR_BAD_DESC;
  if (params == nullptr || (B > 0 && x_in == nullptr)) return GPFLOW_ERR_NULL_POINTER;
  if ((d->flags & GPFLOW_FLAG_EULER_PEN) && zw == nullptr) return GPFLOW_ERR_NULL_POINTER;
  const size_t es = d->dtype == GPFLOW_F64 ? 8 : 4;
  const void* all[] = {params, x_in, x_data, grad_y, reg_cot, zw, y_out, logdet_out, traj_out, grad_x_out, grads_out, scalars_out};
  for (const void* q : all)
    if (misaligned(q, es)) return GPFLOW_ERR_MISALIGNED;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (d->dtype == GPFLOW_F32) {
    Ptrs<float> p = {(const float*)params, (const float*)x_in, (const float*)x_data, (const float*)grad_y,
                     (const float*)reg_cot, (const float*)zw, (float*)y_out, (float*)logdet_out, (float*)traj_out,
                     (float*)grad_x_out, (float*)grads_out, (float*)scalars_out};
    return dispatch<float>(d, op, p, B, ws, wsb, st);
  }
  Ptrs<double> p = {(const double*)params, (const double*)x_in, (const double*)x_data, (const double*)grad_y,
                    (const double*)reg_cot, (const double*)zw, (double*)y_out, (double*)logdet_out, (double*)traj_out,
                    (double*)grad_x_out, (double*)grads_out, (double*)scalars_out};
  return dispatch<double>(d, op, p, B, ws, wsb, st);
}

// ---- FP32 FMA burn (roofline denominator) --------------------------------------------------------
__global__ void __launch_bounds__(256) fma_burn_kernel(long long iters, float* sink) {
  float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f, a4 = a0 + 4.f, a5 = a0 + 5.f,
        a6 = a0 + 6.f, a7 = a0 + 7.f;
  const float m = 0.999f + blockIdx.x * 1e-9f, c = 1e-3f;
  for (long long i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
      a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
    }
  }
  const float s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
  if (s == 123.456f) sink[0] = s;
}

// ---- TF32 mma.sync burn (tensor-pipe roofline denominator of the generic kernel's 3xTF32 contractions) ----
__global__ void __launch_bounds__(512) tf32_mma_burn_kernel(long long iters, float* sink) {
  float c[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f;
  const unsigned a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, b0 = a0 + 4, b1 = a0 + 5;
  for (long long it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                   : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  if (s == 123.456f) sink[0] = s;
}

// ---- SM clock probe: (clock64 ticks, globaltimer ns) over a short spin; MHz = 1e3 * ticks / ns ----
__global__ void clock_probe_kernel(unsigned long long* out) {
  unsigned long long c0 = clock64(), t0, t1;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  while (clock64() - c0 < 40000ull) {
  }
  const unsigned long long c1 = clock64();
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
  out[0] = c1 - c0;
  out[1] = t1 - t0;
}

}  // namespace

// torch.optim.Adam (L2-coupled weight decay) on a flat buffer, element-wise; grid-stride
template <typename T>
__global__ void adam_step_kernel(int64_t n, T* __restrict__ p, const T* __restrict__ g, T* __restrict__ m, T* __restrict__ v,
                                 T step_size, T beta1, T beta2, T eps, T wd, T bc2_sqrt) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const T pi = p[i];
    const T gi = g[i] + wd * pi;
    const T mi = m[i] + (T(1) - beta1) * (gi - m[i]);
    const T vi = beta2 * v[i] + (T(1) - beta2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    p[i] = pi - step_size * (mi / (sqrt(vi) / bc2_sqrt + eps));
  }
}

extern "C" {

int64_t gpflow_param_count(const GpFlowDesc* d) {
  if (validate(d) != GPFLOW_OK) return -1;
  int64_t n = 0, prev = d->d + 1;
  for (int i = 0; i < d->n_hidden; ++i) {
    n += prev * d->hidden[i] + d->hidden[i];
    prev = d->hidden[i];
  }
  n += prev * ((int64_t)d->d * d->nb) + (int64_t)d->d * d->nb;
  return n;
}

int gpflow_query_workspace(const GpFlowDesc* d, int64_t B, int op, size_t* bytes) {
  const int rc = validate(d);
  if (rc != GPFLOW_OK) return rc;
  if (bytes == nullptr) return GPFLOW_ERR_NULL_POINTER;
  if (B < 0 || op < GPFLOW_OP_FORWARD || op > GPFLOW_OP_BACKWARD) return GPFLOW_ERR_BAD_DESC;
  const int64_t np = gpflow_param_count(d);
  SmallShape s;
  if (sf2_match(d, op)) {
    *bytes = sf2_15_15_workspace(d, B, op);
    return GPFLOW_OK;
  }
  if (sf2e_match(d, op)) {
    *bytes = sf2e_15_15_workspace(d, B);
    return GPFLOW_OK;
  }
  if (small_match(d, &s)) {
    *bytes = small_workspace(d, B, op, (int)np);
    return GPFLOW_OK;
  }
  GpFlowDesc d15;
  if (pad2d_match(d, op, &d15)) {
    *bytes = pad2d_workspace(&d15, B, op);
    return GPFLOW_OK;
  }
  if (d->dtype == GPFLOW_F32 && tc_selected(d, B)) {
    *bytes = tc_workspace(d, B, op);
    return GPFLOW_OK;
  }
  *bytes = generic_workspace_bytes(d, B, op);
  return GPFLOW_OK;
}

int gpflow_forward(const GpFlowDesc* d, int64_t B, const void* params, const void* x_in, const void* zw, void* y_out,
                   void* logdet_out, void* traj_out, void* scalars_out, void* ws, size_t wsb, void* stream) {
  if (y_out == nullptr && B > 0) return GPFLOW_ERR_NULL_POINTER;
  return run(d, GPFLOW_OP_FORWARD, params, x_in, nullptr, nullptr, nullptr, zw, y_out, logdet_out, traj_out, nullptr,
             nullptr, scalars_out, B, ws, wsb, stream);
}

int gpflow_train_step(const GpFlowDesc* d, int64_t B, const void* params, const void* x_data, const void* x_map,
                      const void* zw, void* grads_out, void* scalars_out, void* y_out, void* grad_xmap_out, void* ws,
                      size_t wsb, void* stream) {
  if (d != nullptr && !(d->flags & GPFLOW_FLAG_WRAP_GAUSS)) return GPFLOW_ERR_BAD_DESC;
  if (grads_out == nullptr || scalars_out == nullptr || (B > 0 && x_data == nullptr)) return GPFLOW_ERR_NULL_POINTER;
  return run(d, GPFLOW_OP_TRAIN_STEP, params, x_map, x_data, nullptr, nullptr, zw, y_out, nullptr, nullptr,
             grad_xmap_out, grads_out, scalars_out, B, ws, wsb, stream);
}

int gpflow_backward(const GpFlowDesc* d, int64_t B, const void* params, const void* x_in, const void* grad_y,
                    const void* reg_cotangent, const void* zw, void* grads_out, void* grad_x_out, void* ws, size_t wsb,
                    void* stream) {
  if (grads_out == nullptr || (B > 0 && grad_y == nullptr)) return GPFLOW_ERR_NULL_POINTER;
  return run(d, GPFLOW_OP_BACKWARD, params, x_in, nullptr, grad_y, reg_cotangent, zw, nullptr, nullptr, nullptr,
             grad_x_out, grads_out, nullptr, B, ws, wsb, stream);
}

int gpflow_velocity(const GpFlowDesc* d, int64_t B, const void* params, const void* x, double t, void* v_out, void* ws,
                    size_t wsb, void* stream) {
  const int rc = validate(d);
  if (rc != GPFLOW_OK) return rc;
  if (params == nullptr || x == nullptr || v_out == nullptr) return GPFLOW_ERR_NULL_POINTER;
  if (B <= 0) return B == 0 ? GPFLOW_OK : GPFLOW_ERR_BAD_DESC;
  GpFlowDesc dd = *d;
  dd.flags &= ~(uint32_t)GPFLOW_FLAG_EULER_PEN;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int sm = device_info().sm_count;
  if (d->dtype == GPFLOW_F32 && (d->flags & GPFLOW_FLAG_FORCE_TC) && tc_match(&dd))   // workspace: query with FORCE_TC set
    return tc_f32(&dd, GPFLOW_OP_FORWARD, (const float*)params, (const float*)x, nullptr, nullptr, nullptr, nullptr, nullptr,
                  nullptr, nullptr, nullptr, B, ws, wsb, st, sm, (float*)v_out, t);
  if (d->dtype == GPFLOW_F32)
    return generic_f32(&dd, GPFLOW_OP_FORWARD, (const float*)params, (const float*)x, nullptr, nullptr, nullptr,
                                 nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, B, ws, wsb, st, sm,
                                 (float*)v_out, t);
  return generic_f64(&dd, GPFLOW_OP_FORWARD, (const double*)params, (const double*)x, nullptr, nullptr,
                                nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, B, ws, wsb, st,
                                sm, (double*)v_out, t);
}

int gpflow_kernel_variant(const GpFlowDesc* d, int op, char* buf, size_t n) {
  const int rc = validate(d);
  if (rc != GPFLOW_OK) return rc;
  if (buf == nullptr || n == 0) return GPFLOW_ERR_NULL_POINTER;
  const char* ty = d->dtype == GPFLOW_F64 ? "f64" : "f32";
  const char* mode = op == GPFLOW_OP_FORWARD ? "fwd" : (op == GPFLOW_OP_TRAIN_STEP ? "train" : "bwd");
  const char* eu = (d->flags & GPFLOW_FLAG_EULER_PEN) ? ",euler" : "";
  SmallShape s;
  GpFlowDesc d15;
  if (sf2_match(d, op))
    snprintf(buf, n, "%s<f32,d=2,nb=1,H=[15,15],%s>", sf2_15_15_kernel_name(op), mode);
  else if (sf2e_match(d, op))
    snprintf(buf, n, "gpflow_sf2e_kernel<f32,d=2,nb=1,H=[15,15],%s,euler>", mode);
  else if (small_match(d, &s))
    snprintf(buf, n, "gpflow_small_kernel<%s,d=%d,nb=%d,H=[%d,%d],%s%s>", ty, s.d, s.nb, s.h1, s.h2, mode, eu);
  else if (pad2d_match(d, op, &d15))
    snprintf(buf, n, "%s<f32,d=2,nb=1,H=[15,15],%s%s> (H=[%d,%d] zero-padded)",
             sf2_match(&d15, op) ? sf2_15_15_kernel_name(op) : "gpflow_sf2e_kernel", mode, eu, d->hidden[0], d->hidden[1]);
  else if (d->dtype == GPFLOW_F32 && tc_selected(d, (d->flags & GPFLOW_FLAG_FORCE_TC) ? 0 : tc_min_batch(d)))
    snprintf(buf, n, "gpflow_tc<f32,tcgen05,%s> (batches >= %lld; generic below)", mode, (long long)tc_min_batch(d));
  else
    snprintf(buf, n, "gpflow_generic_kernel<%s,%s%s>", ty, mode, eu);
  return GPFLOW_OK;
}

int gpflow_launch_count(const GpFlowDesc* d, int op) {
  if (validate(d) != GPFLOW_OK) return -1;
  (void)op;
  SmallShape s;
  if (sf2_match(d, op)) return sf2_15_15_launches(op);   // [weight pack +] main kernel + deterministic finalize
  if (sf2e_match(d, op)) return 3;
  if (small_match(d, &s)) return 2;            // main kernel + deterministic finalize
  GpFlowDesc d15;
  if (pad2d_match(d, op, &d15))                // + parameter scatter [+ gradient gather]
    return (sf2_match(&d15, op) ? sf2_15_15_launches(op) : 3) + (op == GPFLOW_OP_FORWARD ? 1 : 2);
  return generic_runs_padded(d) ? 4 : 3;       // [zero-padding scatter +] weight transpose + main kernel + finalize
}

int64_t gpflow_tc_launches_total(void) { return (int64_t)tc_launches_total(); }

int gpflow_tc_slab_samples(const GpFlowDesc* d, int64_t B, int op) {
  if (validate(d) != GPFLOW_OK || !tc_match(d)) return -1;
  return tc_slab(d, B, op);
}

int gpflow_fp32_fma_burn(int64_t iters, void* sink, double* host_flops_out, void* stream) {
  if (sink == nullptr || iters < 1) return GPFLOW_ERR_NULL_POINTER;
  const DeviceInfo di = device_info();
  const int grid = di.sm_count * 8, block = 256;
  fma_burn_kernel<<<grid, block, 0, reinterpret_cast<cudaStream_t>(stream)>>>((long long)iters, (float*)sink);
  if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  if (host_flops_out) *host_flops_out = 2.0 * 128.0 * (double)iters * (double)grid * (double)block;
  return GPFLOW_OK;
}

int gpflow_tf32_mma_burn(int64_t iters, void* sink, double* host_flops_out, void* stream) {
  if (sink == nullptr) return GPFLOW_ERR_NULL_POINTER;
  if (iters < 1) return GPFLOW_ERR_BAD_DESC;
  const DeviceInfo di = device_info();
  const int grid = di.sm_count, block = 512;   // 16 warps per SM (measured: saturates from 8)
  tf32_mma_burn_kernel<<<grid, block, 0, reinterpret_cast<cudaStream_t>(stream)>>>((long long)iters, (float*)sink);
  if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  if (host_flops_out) *host_flops_out = 2.0 * 16.0 * 8.0 * 8.0 * 8.0 * (double)iters * (double)grid * (double)(block / 32);
  return GPFLOW_OK;
}

int gpflow_clock_probe(void* out2, void* stream) {
  if (out2 == nullptr) return GPFLOW_ERR_NULL_POINTER;
  clock_probe_kernel<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream)>>>(reinterpret_cast<unsigned long long*>(out2));
  return cudaGetLastError() == cudaSuccess ? GPFLOW_OK : GPFLOW_ERR_CUDA;
}

int gpflow_adam_step(int dtype, int64_t n, void* params, const void* grads, void* exp_avg, void* exp_avg_sq, double lr,
                     double beta1, double beta2, double eps, double weight_decay, int64_t step, void* stream) {
  if (params == nullptr || grads == nullptr || exp_avg == nullptr || exp_avg_sq == nullptr) return GPFLOW_ERR_NULL_POINTER;
  if (n < 0 || step < 1 || (dtype != GPFLOW_F32 && dtype != GPFLOW_F64)) return GPFLOW_ERR_BAD_DESC;
  if (n == 0) return GPFLOW_OK;
  const double bc1 = 1.0 - pow(beta1, (double)step), bc2 = 1.0 - pow(beta2, (double)step);
  const double step_size = lr / bc1, bc2_sqrt = sqrt(bc2);
  const int64_t max_ctas = (int64_t)device_info().sm_count * 8;
  const unsigned grid = (unsigned)((n + 255) / 256 < max_ctas ? (n + 255) / 256 : max_ctas);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == GPFLOW_F32)
    adam_step_kernel<float><<<grid, 256, 0, st>>>(n, (float*)params, (const float*)grads, (float*)exp_avg, (float*)exp_avg_sq,
                                                  (float)step_size, (float)beta1, (float)beta2, (float)eps,
                                                  (float)weight_decay, (float)bc2_sqrt);
  else
    adam_step_kernel<double><<<grid, 256, 0, st>>>(n, (double*)params, (const double*)grads, (double*)exp_avg,
                                                   (double*)exp_avg_sq, step_size, beta1, beta2, eps, weight_decay, bc2_sqrt);
  return cudaGetLastError() == cudaSuccess ? GPFLOW_OK : GPFLOW_ERR_CUDA;
}

static int bnaf_check(const GpFlowBnafDesc* d) {
  if (d == nullptr) return GPFLOW_ERR_NULL_POINTER;
  if (d->abi_version != GPFLOW_ABI_VERSION || d->dim < 1 || d->hidden_per_dim < 1 || d->n_layers < 2) return GPFLOW_ERR_BAD_DESC;
  if (d->dtype != GPFLOW_F32 || d->dim != 2 || d->hidden_per_dim != 16 || d->n_layers > kBnafMaxLayers)
    return GPFLOW_ERR_UNSUPPORTED;
  return GPFLOW_OK;
}

int gpflow_bnaf_param_count(const GpFlowBnafDesc* d, int64_t* count) {
  if (count == nullptr) return GPFLOW_ERR_NULL_POINTER;
  const int rc = bnaf_check(d);
  if (rc != GPFLOW_OK) return rc;
  *count = Bnaf<2, 16>::param_count(d->n_layers);
  return GPFLOW_OK;
}

int gpflow_bnaf_forward(const GpFlowBnafDesc* d, int64_t B, const void* params, const void* x, void* y_out,
                        void* logdet_out, void* stream) {
  const int rc = bnaf_check(d);
  if (rc != GPFLOW_OK) return rc;
  if (B < 0) return GPFLOW_ERR_BAD_DESC;
  if (B == 0) return GPFLOW_OK;
  if (params == nullptr || x == nullptr || y_out == nullptr) return GPFLOW_ERR_NULL_POINTER;
  if (reinterpret_cast<uintptr_t>(params) & 3) return GPFLOW_ERR_MISALIGNED;
  return launch_bnaf_forward<2, 16>(d->n_layers, (const float*)params, (const float*)x, (float*)y_out, (float*)logdet_out,
                                    B, reinterpret_cast<cudaStream_t>(stream), device_info().sm_count);
}

static int ffjord_check(const GpFlowFfjordDesc* d) {
  if (d == nullptr) return GPFLOW_ERR_NULL_POINTER;
  if (d->abi_version != GPFLOW_ABI_VERSION || d->dim < 1 || d->n_hidden < 0 || d->n_hidden > GPFLOW_FFJORD_MAX_HIDDEN ||
      d->n_blocks < 1 || d->n_blocks > GPFLOW_FFJORD_MAX_BLOCKS || d->n_steps < 1 ||
      (d->dtype != GPFLOW_F32 && d->dtype != GPFLOW_F64))
    return GPFLOW_ERR_BAD_DESC;
  for (int l = 0; l < d->n_hidden; ++l)
    if (d->hidden[l] < 1) return GPFLOW_ERR_BAD_DESC;
  return GPFLOW_OK;
}

int gpflow_ffjord_param_count(const GpFlowFfjordDesc* d, int64_t* count) {
  if (count == nullptr) return GPFLOW_ERR_NULL_POINTER;
  const int rc = ffjord_check(d);
  if (rc != GPFLOW_OK) return rc;
  int64_t n = 0;
  int in = d->dim;
  for (int l = 0; l <= d->n_hidden; ++l) {
    const int out = l < d->n_hidden ? d->hidden[l] : d->dim;
    n += (int64_t)in * out + 4 * out;
    in = out;
  }
  *count = n * d->n_blocks;
  return GPFLOW_OK;
}

int gpflow_ffjord_integrate(const GpFlowFfjordDesc* d, int64_t B, const void* params, const void* x, void* y_out,
                            void* logdet_out, void* stream) {
  const int rc = ffjord_check(d);
  if (rc != GPFLOW_OK) return rc;
  if (B < 0) return GPFLOW_ERR_BAD_DESC;
  if (B == 0) return GPFLOW_OK;
  if (params == nullptr || x == nullptr || y_out == nullptr) return GPFLOW_ERR_NULL_POINTER;
  const size_t es = d->dtype == GPFLOW_F64 ? 8 : 4;
  const void* all[] = {params, x, y_out, logdet_out};
  for (const void* q : all)
    if (q != nullptr && (reinterpret_cast<uintptr_t>(q) % es) != 0) return GPFLOW_ERR_MISALIGNED;
  FfjordArgs a;
  a.params = params; a.x = x; a.y = y_out; a.logdet = logdet_out; a.B = B;
  a.d = d->dim; a.n_linear = d->n_hidden + 1; a.n_blocks = d->n_blocks; a.n_steps = d->n_steps; a.reverse = d->reverse != 0;
  a.width[0] = d->dim;
  for (int l = 0; l < d->n_hidden; ++l) a.width[l + 1] = d->hidden[l];
  a.width[d->n_hidden + 1] = d->dim;
  for (int l = d->n_hidden + 2; l <= kFfjMaxLinear; ++l) a.width[l] = 0;
  for (int b = 0; b < kFfjMaxBlocks; ++b) a.t_end[b] = b < d->n_blocks ? d->t_end[b] : 0.0;
  const size_t smem_limit = 227 * 1024;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  return d->dtype == GPFLOW_F64 ? launch_ffjord<double>(a, st, device_info().sm_count, smem_limit)
                                : launch_ffjord<float>(a, st, device_info().sm_count, smem_limit);
}

const char* gpflow_error_string(int status) {
  switch (status) {
    case GPFLOW_OK: return "ok";
    case GPFLOW_ERR_BAD_DESC: return "bad descriptor (inconsistent or out-of-range field)";
    case GPFLOW_ERR_UNSUPPORTED: return "configuration not supported by the CUDA kernels";
    case GPFLOW_ERR_NULL_POINTER: return "required pointer is NULL";
    case GPFLOW_ERR_WORKSPACE: return "workspace too small (see gpflow_query_workspace)";
    case GPFLOW_ERR_MISALIGNED: return "buffer not aligned";
    case GPFLOW_ERR_CUDA: return "CUDA launch failed";
    case GPFLOW_ERR_STATIC_SCRATCH:
      return "gpflow_velocity_static: batch / shape needs more than GPFLOW_STATIC_SCRATCH_BYTES of scratch "
             "(use gpflow_velocity with a caller-provided workspace)";
    case GPFLOW_ERR_COLLECTIVE:
      return "multi-GPU hook: NCCL could not be loaded at run time (libnccl.so.2 / $GPFLOW_NCCL_LIB) or returned an error";
    default: return "unknown status";
  }
}

int gpflow_abi_version(void) { return GPFLOW_ABI_VERSION; }

}  // extern "C"
