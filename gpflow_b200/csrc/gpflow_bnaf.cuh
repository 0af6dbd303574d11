// gpflow_b200 -- forward of the frozen BNAF map that feeds the GP flow, fp32.
//
// `x_map = self.map(x)[0]` under no_grad is the step immediately in front of the hot path (src/gp_flows/network.py:76-78);
// for the toy configs (C1 / C2) the map is a block neural autoregressive flow, flow_models/bnaf/bnaf.py:11-39 with one
// block: MaskedWeight(dim -> H) Tanh [MaskedWeight(H -> H) Tanh]* MaskedWeight(H -> dim), H = dim * A.  One launch
// computes y and log|det J| for the whole batch (the eager module runs ~40 small kernels and materialises a
// [B, dim, A, A] tensor per layer for the logsumexp):
//   * MaskedWeight.get_weights (bnaf_module.py:229-249): w = exp(W) mask_d + W mask_o, weight-normalised rows
//     w_eff = exp(dw) w / |w|, log diagonal blocks wpl = dw + W - log|w| -- parameter-only work, redone per CTA into
//     shared memory (2372 parameters for the pretrained maps);
//   * MaskedWeight.forward (:251-277): a = h w_eff^T + b;  g'[db][oo] = logsumexp_kk (wpl[db][oo][kk] + g[db][kk]);
//   * Tanh.forward (:301-325): h = tanh(a), g += -2 (a - log 2 + softplus(-2a));
//   * BNAFModule.forward, res=None (:67-97): log|det| = sum_db g_last[db].
// One THREAD per sample; the sample's (h, g) row lives in shared memory (68-float stride, odd in 16-byte units) and is
// updated IN PLACE from the last diagonal block to the first: the lower-block-triangular mask makes output block db
// depend on input blocks <= db only.  Loops over the input unit are rolled (the weights are warp-uniform LDS.128
// broadcasts), the A outputs of a block are register accumulators.
#pragma once
#include "../../include/gpflow_b200.h"
#include "gpflow_common.cuh"

namespace gpflow {

constexpr int kBnafThreads = 256;
constexpr int kBnafMaxLayers = 8;   // masked linear layers

struct BnafArgs {
  const float* params;   // per layer: _weight [out][in], _diag_weight [out], bias [out]   (state_dict order)
  const float* x;        // [B, DIM]
  float* y;              // [B, DIM]
  float* logdet;         // [B] nullable
  long long B;
  int n_layers;          // masked linear layers (>= 2)
};

template <int DIM, int A>
struct Bnaf {
  static constexpr int H = DIM * A;
  static constexpr int ROW = 2 * H + 4;            // h[H] g[H] + pad: (2H+4)/4 odd for H = 32
  // shared-memory weights of one layer (floats)
  static constexpr int first_size = H * DIM + H + H;          // wT[k<DIM][H], wpl[H], bias[H]
  static constexpr int mid_size = H * H + DIM * A * A + H;    // wT[k][o], wplT[db][kk][oo], bias[H]
  static constexpr int last_size = DIM * H + DIM * A + 4;     // w[db][k], wpl[db][kk], bias[DIM]
  __host__ __device__ static constexpr int weights_floats(int n_layers) {
    return first_size + (n_layers - 2) * mid_size + last_size;
  }
  __host__ __device__ static constexpr int param_count(int n_layers) {
    return (H * DIM + 2 * H) + (n_layers - 2) * (H * H + 2 * H) + (DIM * H + 2 * DIM);
  }
};

GPF_DEV float bnaf_log_dtanh(float a) {  // log(1 - tanh(a)^2) = -2 (a - log 2 + softplus(-2a)), bnaf_module.py:318-320
  const float z = -2.0f * a;
  const float sp = fmaxf(z, 0.0f) + log1pf(__expf(-fabsf(z)));
  return -2.0f * (a - 0.6931471805599453f + sp);
}

// weight normalisation of one MaskedWeight into shared memory; `raw` = [_weight | _diag_weight | bias] in global memory.
// TRANSPOSE: store w_eff as wT[k][o] (k-major) and wpl as wplT[db][kk][oo]; else w[o][k], wpl[o][kk].
template <int DIM>
__device__ void bnaf_normalise_layer(const float* __restrict__ raw, int out, int in, bool transpose, float* __restrict__ w_s,
                                     float* __restrict__ wpl_s, float* __restrict__ bias_s) {
  const int ob = out / DIM, ib = in / DIM;
  const float* W = raw;
  const float* dw = raw + out * in;
  const float* bias = dw + out;
  for (int o = threadIdx.x; o < out; o += blockDim.x) {
    const int db = o / ob, oo = o - db * ob;
    float n2 = 0.0f;
    for (int k = 0; k < (db + 1) * ib; ++k) {
      const float wr = W[o * in + k];
      const float w = k >= db * ib ? expf(wr) : wr;
      n2 = fmaf(w, w, n2);
    }
    const float scale = expf(dw[o]) / sqrtf(n2), lhalf = 0.5f * logf(n2);
    for (int k = 0; k < in; ++k) {
      const float wr = W[o * in + k];
      float w = 0.0f;
      if (k < (db + 1) * ib) w = (k >= db * ib ? expf(wr) : wr) * scale;
      w_s[transpose ? k * out + o : o * in + k] = w;
    }
    for (int kk = 0; kk < ib; ++kk) {
      const float v = dw[o] + W[o * in + db * ib + kk] - lhalf;
      wpl_s[transpose ? (db * ib + kk) * ob + oo : o * ib + kk] = v;
    }
    bias_s[o] = bias[o];
  }
}

template <int DIM, int A>
__global__ void __launch_bounds__(kBnafThreads, 2) gpflow_bnaf_forward_kernel(const BnafArgs a) {
  typedef Bnaf<DIM, A> K;
  constexpr int H = K::H;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* const Wsm = reinterpret_cast<float*>(smem_raw);
  float* const rows = Wsm + round_up(K::weights_floats(a.n_layers), 4);
  const int tid = threadIdx.x;

  // ---- parameter-only prologue: MaskedWeight.get_weights of every layer
  {
    const float* raw = a.params;
    float* s = Wsm;
    bnaf_normalise_layer<DIM>(raw, H, DIM, true, s, s + H * DIM, s + H * DIM + H);   // first: ib = 1, wpl[o]
    raw += H * DIM + 2 * H;
    s += K::first_size;
    for (int l = 0; l < a.n_layers - 2; ++l) {
      bnaf_normalise_layer<DIM>(raw, H, H, true, s, s + H * H, s + H * H + DIM * A * A);
      raw += H * H + 2 * H;
      s += K::mid_size;
    }
    bnaf_normalise_layer<DIM>(raw, DIM, H, false, s, s + DIM * H, s + DIM * H + DIM * A);
  }
  __syncthreads();

  float* const hrow = rows + (size_t)tid * K::ROW;
  float* const grow = hrow + H;
  for (long long b = (long long)blockIdx.x * blockDim.x + tid; b < a.B; b += (long long)gridDim.x * blockDim.x) {
    float xin[DIM];
#pragma unroll
    for (int j = 0; j < DIM; ++j) xin[j] = a.x[b * DIM + j];
    // ---- first layer (in = DIM, one input per block) + Tanh
    {
      const float* wT = Wsm;                // [k][o]
      const float* wpl = Wsm + H * DIM;     // [o]
      const float* bias = wpl + H;
#pragma unroll 1
      for (int o = 0; o < H; o += 4) {
        float4 acc = *reinterpret_cast<const float4*>(bias + o);
#pragma unroll
        for (int k = 0; k < DIM; ++k) {     // masked entries of w_eff are stored as zeros
          const float4 w = *reinterpret_cast<const float4*>(wT + k * H + o);
          acc.x = fmaf(xin[k], w.x, acc.x); acc.y = fmaf(xin[k], w.y, acc.y);
          acc.z = fmaf(xin[k], w.z, acc.z); acc.w = fmaf(xin[k], w.w, acc.w);
        }
        const float4 g0 = *reinterpret_cast<const float4*>(wpl + o);
        *reinterpret_cast<float4*>(hrow + o) = make_float4(tanh_real(acc.x), tanh_real(acc.y), tanh_real(acc.z), tanh_real(acc.w));
        *reinterpret_cast<float4*>(grow + o) = make_float4(g0.x + bnaf_log_dtanh(acc.x), g0.y + bnaf_log_dtanh(acc.y),
                                                          g0.z + bnaf_log_dtanh(acc.z), g0.w + bnaf_log_dtanh(acc.w));
      }
    }
    // ---- hidden layers + Tanh, in place from the last diagonal block to the first
    const float* ws = Wsm + K::first_size;
#pragma unroll 1
    for (int l = 0; l < a.n_layers - 2; ++l, ws += K::mid_size) {
      const float* wT = ws;                       // [k][o]
      const float* wplT = ws + H * H;             // [db][kk][oo]
      const float* bias = wplT + DIM * A * A;
#pragma unroll 1
      for (int db = DIM - 1; db >= 0; --db) {
        float acc[A], m[A], s[A];
#pragma unroll
        for (int oo = 0; oo < A; ++oo) {
          acc[oo] = bias[db * A + oo];
          m[oo] = -INFINITY;
          s[oo] = 0.0f;
        }
#pragma unroll 2
        for (int k = 0; k < (db + 1) * A; ++k) {
          const float hk = hrow[k];
#pragma unroll
          for (int q = 0; q < A; q += 4) {
            const float4 w = *reinterpret_cast<const float4*>(wT + k * H + db * A + q);
            acc[q] = fmaf(hk, w.x, acc[q]); acc[q + 1] = fmaf(hk, w.y, acc[q + 1]);
            acc[q + 2] = fmaf(hk, w.z, acc[q + 2]); acc[q + 3] = fmaf(hk, w.w, acc[q + 3]);
          }
        }
        // logsumexp over the inputs of the diagonal block: max pass, then sum pass (torch.logsumexp)
#pragma unroll 2
        for (int kk = 0; kk < A; ++kk) {
          const float gk = grow[db * A + kk];
#pragma unroll
          for (int q = 0; q < A; q += 4) {
            const float4 w = *reinterpret_cast<const float4*>(wplT + (db * A + kk) * A + q);
            m[q] = fmaxf(m[q], w.x + gk); m[q + 1] = fmaxf(m[q + 1], w.y + gk);
            m[q + 2] = fmaxf(m[q + 2], w.z + gk); m[q + 3] = fmaxf(m[q + 3], w.w + gk);
          }
        }
#pragma unroll 2
        for (int kk = 0; kk < A; ++kk) {
          const float gk = grow[db * A + kk];
#pragma unroll
          for (int q = 0; q < A; q += 4) {
            const float4 w = *reinterpret_cast<const float4*>(wplT + (db * A + kk) * A + q);
            s[q] += __expf(w.x + gk - m[q]); s[q + 1] += __expf(w.y + gk - m[q + 1]);
            s[q + 2] += __expf(w.z + gk - m[q + 2]); s[q + 3] += __expf(w.w + gk - m[q + 3]);
          }
        }
#pragma unroll
        for (int oo = 0; oo < A; ++oo) {
          hrow[db * A + oo] = tanh_real(acc[oo]);
          grow[db * A + oo] = m[oo] + logf(s[oo]) + bnaf_log_dtanh(acc[oo]);
        }
      }
    }
    // ---- last layer (out = DIM): y and the log-determinant
    {
      const float* w = ws;                 // [db][k]
      const float* wpl = ws + DIM * H;     // [db][kk]
      const float* bias = wpl + DIM * A;
      float ld = 0.0f;
#pragma unroll
      for (int db = 0; db < DIM; ++db) {
        float acc = bias[db], m = -INFINITY, s = 0.0f;
#pragma unroll 4
        for (int k = 0; k < (db + 1) * A; ++k) acc = fmaf(hrow[k], w[db * H + k], acc);
#pragma unroll 4
        for (int kk = 0; kk < A; ++kk) m = fmaxf(m, wpl[db * A + kk] + grow[db * A + kk]);
#pragma unroll 4
        for (int kk = 0; kk < A; ++kk) s += __expf(wpl[db * A + kk] + grow[db * A + kk] - m);
        a.y[b * DIM + db] = acc;
        ld += m + logf(s);
      }
      if (a.logdet != nullptr) a.logdet[b] = ld;
    }
  }
}

template <int DIM, int A>
int launch_bnaf_forward(int n_layers, const float* params, const float* x, float* y, float* logdet, int64_t B,
                        cudaStream_t st, int sm_count) {
  typedef Bnaf<DIM, A> K;
  if (n_layers < 2 || n_layers > kBnafMaxLayers) return GPFLOW_ERR_UNSUPPORTED;
  auto kern = gpflow_bnaf_forward_kernel<DIM, A>;
  const size_t smem = (size_t)(round_up(K::weights_floats(n_layers), 4) + kBnafThreads * K::ROW) * sizeof(float);
  if (smem > 113 * 1024) return GPFLOW_ERR_UNSUPPORTED;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
    cudaGetLastError();
    return GPFLOW_ERR_CUDA;
  }
  BnafArgs a;
  a.params = params; a.x = x; a.y = y; a.logdet = logdet; a.B = B; a.n_layers = n_layers;
  int64_t grid = (B + kBnafThreads - 1) / kBnafThreads;
  if (grid > 2 * sm_count) grid = 2 * sm_count;   // persistent: 2 CTAs of 8 warps per SM
  if (grid < 1) grid = 1;
  kern<<<(unsigned)grid, kBnafThreads, smem, st>>>(a);
  return cudaGetLastError() == cudaSuccess ? GPFLOW_OK : GPFLOW_ERR_CUDA;
}

}  // namespace gpflow
