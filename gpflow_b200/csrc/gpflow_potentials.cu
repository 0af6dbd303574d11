// gpflow_b200 -- standalone VectorField.eval / VectorField.jacobian (src/gp_flows/vector_field.py:75-128): the tanh MLP
// potentials f(x) and their Jacobian df/dx by forward-mode propagation, for ALL inputs of the network (the velocity field
// feeds it [x; t]: the Jacobian returned here has the time column too, like the reference's).
// Not on the hot path (the fused kernels evaluate both inside the velocity); this op exists so that the mirror's
// `VectorField.eval / .jacobian / __call__` work on their own (plots, the reference's tests/test_vector_field.py).
// One CTA per sample at a time (grid-stride), activations and tangents of two consecutive layers in shared memory.
#include "../../include/gpflow_b200.h"

#include <cuda_runtime.h>
#include <stdint.h>

#include "gpflow_common.cuh"

namespace gpflow {

struct PotArgs {
  int n_lin;                            // number of nn.Linear layers
  int width[GPFLOW_MAX_HIDDEN + 2];     // dim_in, hidden..., dim_out
  int woff[GPFLOW_MAX_HIDDEN + 1], boff[GPFLOW_MAX_HIDDEN + 1];
  int hmax;                             // widest hidden layer
  long long B;
};

template <typename T>
__global__ void __launch_bounds__(128) gpflow_potentials_kernel(const PotArgs a, const T* __restrict__ params,
                                                                const T* __restrict__ x, T* __restrict__ f_out,
                                                                T* __restrict__ jac_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int din = a.width[0], dout = a.width[a.n_lin];
  const int row = din + 1;                                   // [h | T[0..din)] per unit
  T* buf0 = reinterpret_cast<T*>(smem_raw);
  T* buf1 = buf0 + (size_t)a.hmax * row;
  T* xin = buf1 + (size_t)a.hmax * row;                      // [din]
  for (long long b = blockIdx.x; b < a.B; b += gridDim.x) {
    for (int j = threadIdx.x; j < din; j += blockDim.x) xin[j] = x[b * din + j];
    __syncthreads();
    T* cur = buf0;
    T* nxt = buf1;
    for (int l = 0; l < a.n_lin; ++l) {
      const int nin = a.width[l], nout = a.width[l + 1];
      const bool last = l == a.n_lin - 1;
      const T* W = params + a.woff[l];
      const T* bias = params + a.boff[l];
      // one (unit, column) pair per thread and iteration; column 0 is the value, columns 1..din the tangents
      for (int idx = threadIdx.x; idx < nout * row; idx += blockDim.x) {
        const int k = idx / row, c = idx % row;
        T s;
        if (l == 0) {
          if (c == 0) {
            s = bias[k];
            for (int m = 0; m < nin; ++m) s += W[k * nin + m] * xin[m];
          } else {
            s = W[k * nin + (c - 1)];
          }
        } else {
          s = c == 0 ? bias[k] : T(0);
          for (int m = 0; m < nin; ++m) s += W[k * nin + m] * cur[m * row + c];
        }
        if (last) {
          if (c == 0) {
            if (f_out != nullptr) f_out[b * dout + k] = s;
          } else if (jac_out != nullptr) {
            jac_out[(b * dout + k) * din + (c - 1)] = s;
          }
        } else {
          nxt[idx] = s;   // pre-activation value / un-scaled tangent
        }
      }
      __syncthreads();
      if (!last) {   // h = tanh(a), T = (1 - h^2) U   (vector_field.py:84-88,119-126)
        for (int k = threadIdx.x; k < nout; k += blockDim.x) {
          const T h = act_tanh(nxt[k * row]);
          const T sp = T(1) - h * h;
          nxt[k * row] = h;
          for (int c = 1; c < row; ++c) nxt[k * row + c] *= sp;
        }
        __syncthreads();
        T* t = cur; cur = nxt; nxt = t;
      }
    }
    __syncthreads();
  }
}

}  // namespace gpflow

using namespace gpflow;

extern "C" int gpflow_potentials(int dtype, int n_linear, const int32_t* widths, int64_t B, const void* params, const void* x,
                                 void* f_out, void* jac_out, void* cuda_stream) {
  if (widths == nullptr || params == nullptr || (B > 0 && x == nullptr)) return GPFLOW_ERR_NULL_POINTER;
  if (dtype != GPFLOW_F32 && dtype != GPFLOW_F64) return GPFLOW_ERR_BAD_DESC;
  if (n_linear < 1 || n_linear > GPFLOW_MAX_HIDDEN + 1 || B < 0) return GPFLOW_ERR_BAD_DESC;
  if (B == 0) return GPFLOW_OK;
  PotArgs a;
  a.n_lin = n_linear;
  a.B = B;
  a.hmax = 1;
  int off = 0;
  for (int l = 0; l <= n_linear; ++l) {
    if (widths[l] < 1) return GPFLOW_ERR_BAD_DESC;
    a.width[l] = widths[l];
    if (l >= 1 && l < n_linear && widths[l] > a.hmax) a.hmax = widths[l];
  }
  for (int l = 0; l < n_linear; ++l) {   // flat layout [W_1, b_1, ..., W_n, b_n], W row-major [out, in] (vector_field.py:56-61)
    a.woff[l] = off;
    off += a.width[l] * a.width[l + 1];
    a.boff[l] = off;
    off += a.width[l + 1];
  }
  const size_t es = dtype == GPFLOW_F64 ? 8 : 4;
  const size_t smem = ((size_t)2 * a.hmax * (a.width[0] + 1) + a.width[0]) * es + 16;
  if (smem > 200 * 1024) return GPFLOW_ERR_UNSUPPORTED;   // e.g. d = 64, H = 256 in fp64; the fused kernels have no such limit
  cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
  const unsigned grid = (unsigned)(B < 148 * 8 ? B : 148 * 8);
  cudaError_t e;
  if (dtype == GPFLOW_F32) {
    e = cudaFuncSetAttribute(gpflow_potentials_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess)
      gpflow_potentials_kernel<float><<<grid, 128, smem, st>>>(a, (const float*)params, (const float*)x, (float*)f_out,
                                                              (float*)jac_out);
  } else {
    e = cudaFuncSetAttribute(gpflow_potentials_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess)
      gpflow_potentials_kernel<double><<<grid, 128, smem, st>>>(a, (const double*)params, (const double*)x, (double*)f_out,
                                                               (double*)jac_out);
  }
  if (e != cudaSuccess || cudaGetLastError() != cudaSuccess) {
    cudaGetLastError();
    return GPFLOW_ERR_CUDA;
  }
  return GPFLOW_OK;
}
