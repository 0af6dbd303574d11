// gpflow_b200 -- the frozen FFJORD map either side of the GP flow, fp32 / fp64.
//
// `map.backward(gauss)` builds the Gaussian-noise training set of the latent configs (src/data_module.py:161-186,
// `FFJORD.backward`, flow_models/ffjord/ffjord.py:55-56) and `x_map, log_det = self.map(x)` is the step in front of the
// train kernel (src/gp_flows/network.py:76-78, ffjord.py:50-53).  The map is a chain of CNF blocks
// (layers/container.py:14-26) whose vector field is an `ODEnet` of ConcatSquashLinear layers with tanh in between
// (layers/odefunc.py:160-174, layers/diffeq_layers/basic.py:76-86):
//        a_l = (W_l h_{l-1} + b_l) * sigmoid(gw_l t + gb_l) + hb_l t,     h_l = tanh(a_l)  (none after the last layer)
// and whose log-density state integrates minus the exact trace of df/dy in eval mode (`divergence_bf`, odefunc.py:13-17,
// :268-301).  The reference integrates with torchdiffeq's adaptive dopri5 (cnf.py:66-83; not vendored): this kernel
// integrates every block with `n_steps` classical RK4 steps (the GP flow's own scheme) -- see oracle/ffjord_oracle.py.
//
// One launch for the whole chain and the whole batch.  One THREAD per sample; its state, stage vectors, hidden
// activations and tangent columns live in shared memory in column layout ([k][thread], conflict-free), the weights of
// the current block are staged transposed ([in][out], zero-padded to 8 outputs) so that a k-step is one activation
// load + two warp-uniform 128-bit weight loads feeding 8 accumulators.  The time-only factors (gates, hyper-bias) are
// recomputed by the CTA once per distinct stage time.  The trace is propagated one input direction at a time through
// the stored activations (d extra hidden matvecs per evaluation) -- only when log_det is requested.
#pragma once
#include "../../include/gpflow_b200.h"
#include "gpflow_common.cuh"

namespace gpflow {

constexpr int kFfjMaxLinear = 8;   // linear layers per ODEnet (hidden layers + 1)
constexpr int kFfjMaxBlocks = 8;
constexpr int kFfjOB = 8;          // outputs per register block
constexpr int kFfjMaxThreads = 256;

struct FfjordArgs {
  const void* params;
  const void* x;
  void* y;
  void* logdet;      // nullable
  long long B;
  int d, n_linear, n_blocks, n_steps, reverse;
  int width[kFfjMaxLinear + 1];    // n_0 = d, hidden widths, n_last = d
  double t_end[kFfjMaxBlocks];
};

// ---- shared-memory plan (element counts), identical on host and device ----
struct FfjordPlan {
  int w_off[kFfjMaxLinear];      // transposed weights [in][outp]
  int v_off[kFfjMaxLinear];      // 6 vectors of outp: b, hb, gw, gb, gate, hbt
  int outp[kFfjMaxLinear];
  int weights_total;             // elements of the weight region
  int dp, hmaxp;
  int act_off[kFfjMaxLinear];    // per-thread column index of hidden activation l (1..n_linear-1)
  int col_y, col_ys, col_acc, col_k, col_ta, col_tb;
  int cols;                      // per-thread columns
};

__host__ __device__ inline FfjordPlan ffjord_plan(const int* width, int n_linear, bool want_logdet) {
  FfjordPlan p;
  int o = 0;
  p.hmaxp = 0;
  for (int l = 0; l < n_linear; ++l) {
    p.outp[l] = round_up(width[l + 1], kFfjOB);
    p.w_off[l] = o; o += width[l] * p.outp[l];
    p.v_off[l] = o; o += 6 * p.outp[l];
    if (l < n_linear - 1 && p.outp[l] > p.hmaxp) p.hmaxp = p.outp[l];
  }
  p.weights_total = round_up(o, 4);
  p.dp = round_up(width[0], kFfjOB);
  int c = 0;
  p.col_y = c; c += p.dp;
  p.col_ys = c; c += p.dp;
  p.col_acc = c; c += p.dp;
  p.col_k = c; c += p.dp;
  for (int l = 0; l < n_linear - 1; ++l) { p.act_off[l] = c; c += p.outp[l]; }
  p.col_ta = p.col_tb = c;
  if (want_logdet) { p.col_ta = c; c += p.hmaxp; p.col_tb = c; c += p.hmaxp; }
  p.cols = c;
  return p;
}

template <int OB> GPF_DEV void ffj_loadw(const float* p, float (&w)[OB]) {   // OB consecutive weights, 16-byte aligned
#pragma unroll
  for (int i = 0; i < OB / 4; ++i) {
    const float4 a = *reinterpret_cast<const float4*>(p + 4 * i);
    w[4 * i] = a.x; w[4 * i + 1] = a.y; w[4 * i + 2] = a.z; w[4 * i + 3] = a.w;
  }
}
template <int OB> GPF_DEV void ffj_loadw(const double* p, double (&w)[OB]) {
#pragma unroll
  for (int i = 0; i < OB / 2; ++i) {
    const double2 a = *reinterpret_cast<const double2*>(p + 2 * i);
    w[2 * i] = a.x; w[2 * i + 1] = a.y;
  }
}
// tanh for the ODEnet activations: 1 - 2 / (2^(2 x log2 e) + 1) on MUFU.EX2 / MUFU.RCP, one regime (absolute error
// ~1.5e-7: the flow's state is O(1) and the test bar is 1e-5 norm-wise, so the relative accuracy near 0 that
// gpflow_common.cuh's two-regime tanh_real pays ~14 more instructions for is not needed here).  fp64: libm.
GPF_DEV float ffj_tanh(float x) {
  const float e = ex2_approx(x * 2.885390043258667f);
  return fmaf(-2.0f, rcp_approx(e + 1.0f), 1.0f);
}
GPF_DEV double ffj_tanh(double x) { return tanh(x); }
GPF_DEV float ffj_sigmoid(float z) { return 1.0f / (1.0f + expf(-z)); }
GPF_DEV double ffj_sigmoid(double z) { return 1.0 / (1.0 + exp(-z)); }

// out_col[o] = act( (W in + b)[o] * gate[o] + hbt[o] ), o < outp (padded outputs come out as act(0)); TANGENT gives
// the tangent form  out_col[o] = (1 - h[o]^2) * gate[o] * (W in)[o]  with h = `h_col` (hidden layers only).
// A thread owns S samples (columns `q * NTH` apart, column stride VT = S * NTH) and OB outputs per register block, so a
// k-step is S activation loads + OB/4 warp-uniform 128-bit weight loads (2 shared-memory wavefronts each) for S * OB
// FMAs: with S = 1 the LSU data pipe caps the FMA pipe at ~45 % (ncu: 66 % LSU wavefronts at 28 % FMA), S = 2 halves
// the weight traffic per FMA.  The operands of k-step k+1 are fetched before the FMAs of k-step k issue (register double
// buffer); addresses advance by pointer increments.
template <typename T, int S, int OB, bool TANGENT>
GPF_DEV void ffj_matvec_ob(const T* __restrict__ Wt, const T* __restrict__ vec, int in, int outp, const T* in_col,
                           T* out_col, const T* h_col, bool hidden, int NTH, int VT) {
  const T* b = vec;
  const T* gate = vec + 4 * outp;
  const T* hbt = vec + 5 * outp;
  for (int o0 = 0; o0 < outp; o0 += OB) {
    T acc[S][OB];
#pragma unroll
    for (int q = 0; q < S; ++q)
#pragma unroll
      for (int i = 0; i < OB; ++i) acc[q][i] = TANGENT ? T(0) : b[o0 + i];
    const T* ip = in_col;
    const T* wq = Wt + o0;
    T hc[S], wc[OB];
#pragma unroll
    for (int q = 0; q < S; ++q) hc[q] = ip[q * NTH];
    ffj_loadw<OB>(wq, wc);
#pragma unroll 2
    for (int k = 1; k < in; ++k) {
      ip += VT;
      wq += outp;
      T hn[S], wn[OB];
#pragma unroll
      for (int q = 0; q < S; ++q) hn[q] = ip[q * NTH];
      ffj_loadw<OB>(wq, wn);
#pragma unroll
      for (int q = 0; q < S; ++q)
#pragma unroll
        for (int i = 0; i < OB; ++i) acc[q][i] = fma(wc[i], hc[q], acc[q][i]);
#pragma unroll
      for (int q = 0; q < S; ++q) hc[q] = hn[q];
#pragma unroll
      for (int i = 0; i < OB; ++i) wc[i] = wn[i];
    }
#pragma unroll
    for (int q = 0; q < S; ++q)
#pragma unroll
      for (int i = 0; i < OB; ++i) acc[q][i] = fma(wc[i], hc[q], acc[q][i]);
#pragma unroll
    for (int i = 0; i < OB; ++i) {
      const T gt = gate[o0 + i], hb = TANGENT ? T(0) : hbt[o0 + i];
#pragma unroll
      for (int q = 0; q < S; ++q) {
        T r;
        if (TANGENT) {
          const T h = h_col[(o0 + i) * VT + q * NTH];
          r = (T(1) - h * h) * (gt * acc[q][i]);
        } else {
          r = acc[q][i] * gt + hb;
          if (hidden) r = ffj_tanh(r);
        }
        out_col[(o0 + i) * VT + q * NTH] = r;
      }
    }
  }
}
template <typename T, int S, bool TANGENT>
GPF_DEV void ffj_matvec(const T* __restrict__ Wt, const T* __restrict__ vec, int in, int outp, const T* in_col,
                        T* out_col, const T* h_col, bool hidden, int NTH, int VT) {
  if (sizeof(T) == 4 && (outp & 15) == 0)
    ffj_matvec_ob<T, S, 16, TANGENT>(Wt, vec, in, outp, in_col, out_col, h_col, hidden, NTH, VT);
  else
    ffj_matvec_ob<T, S, kFfjOB, TANGENT>(Wt, vec, in, outp, in_col, out_col, h_col, hidden, NTH, VT);
}

template <typename T, int S>
__global__ void __launch_bounds__(kFfjMaxThreads, 1) gpflow_ffjord_kernel(const FfjordArgs a) {
  extern __shared__ __align__(16) unsigned char ffj_smem_raw[];
  T* sw = reinterpret_cast<T*>(ffj_smem_raw);
  const bool want_ld = a.logdet != nullptr;
  const FfjordPlan P = ffjord_plan(a.width, a.n_linear, want_ld);
  const int NTH = blockDim.x, tid = threadIdx.x;
  const int VT = S * NTH;                   // samples per CTA tile = column stride
  T* cols = sw + P.weights_total + tid;     // column c of this thread's sample q: cols[c * VT + q * NTH]
  const int d = a.d, nl = a.n_linear;
  const T* gx = static_cast<const T*>(a.x);
  T* gy = static_cast<T*>(a.y);
  T* gld = static_cast<T*>(a.logdet);
  const T* gparams = static_cast<const T*>(a.params);
  long long block_params = 0;
  for (int l = 0; l < nl; ++l) block_params += (long long)a.width[l] * a.width[l + 1] + 4 * a.width[l + 1];

  const long long n_tiles = (a.B + VT - 1) / VT;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    // state
    T logp[S];
#pragma unroll
    for (int q = 0; q < S; ++q) {
      const long long s = tile * VT + q * NTH + tid;
      for (int j = 0; j < P.dp; ++j) cols[(P.col_y + j) * VT + q * NTH] = (s < a.B && j < d) ? gx[s * d + j] : T(0);
      logp[q] = T(0);
    }
    for (int bb = 0; bb < a.n_blocks; ++bb) {
      const int bi = a.reverse ? a.n_blocks - 1 - bb : bb;
      // ---- stage the block's weights: transposed + zero-padded, vectors b | hb | gw | gb ----
      __syncthreads();
      {
        const T* src = gparams + (long long)bi * block_params;
        for (int l = 0; l < nl; ++l) {
          const int in = a.width[l], out = a.width[l + 1], outp = P.outp[l];
          T* Wt = sw + P.w_off[l];
          T* vec = sw + P.v_off[l];
          for (int e = tid; e < in * outp; e += NTH) {
            const int k = e / outp, o = e - k * outp;
            Wt[e] = o < out ? src[o * in + k] : T(0);
          }
          for (int e = tid; e < 4 * outp; e += NTH) {
            const int v = e / outp, o = e - v * outp;
            vec[e] = o < out ? src[in * out + v * out + o] : T(0);
          }
          src += in * out + 4 * out;
        }
      }
      const double T_end = a.t_end[bi];
      const double t0 = a.reverse ? T_end : 0.0, t1 = a.reverse ? 0.0 : T_end;
      const double hd = (t1 - t0) / a.n_steps;
      const T h = (T)hd, hh = (T)(0.5 * hd), h6 = (T)(hd / 6.0);
      for (int step = 0; step < a.n_steps; ++step) {
        const double tk = t0 + step * hd;
#pragma unroll
        for (int q = 0; q < S; ++q)
          for (int j = 0; j < P.dp; ++j) cols[(P.col_acc + j) * VT + q * NTH] = T(0);
        T dacc[S];
#pragma unroll
        for (int q = 0; q < S; ++q) dacc[q] = T(0);
#pragma unroll 1
        for (int st = 0; st < 4; ++st) {
          if (st != 2) {   // stages 1 and 2 share t + h/2
            const double ts = st == 0 ? tk : (st == 1 ? tk + 0.5 * hd : tk + hd);
            const T tt = (T)ts;
            __syncthreads();
            for (int l = 0; l < nl; ++l) {
              const int outp = P.outp[l];
              T* vec = sw + P.v_off[l];
              for (int o = tid; o < outp; o += NTH) {
                vec[4 * outp + o] = ffj_sigmoid(vec[2 * outp + o] * tt + vec[3 * outp + o]);
                vec[5 * outp + o] = vec[outp + o] * tt;
              }
            }
            __syncthreads();
          }
          // stage input
          const T cs = st == 0 ? T(0) : (st == 3 ? h : hh);
#pragma unroll
          for (int q = 0; q < S; ++q) {
            T* c = cols + q * NTH;
            for (int j = 0; j < d; ++j)
              c[(P.col_ys + j) * VT] = st == 0 ? c[(P.col_y + j) * VT] : c[(P.col_y + j) * VT] + cs * c[(P.col_k + j) * VT];
          }
          // value pass
          const T* in_col = cols + P.col_ys * VT;
          for (int l = 0; l < nl; ++l) {
            const bool hidden = l < nl - 1;
            T* out_col = cols + (hidden ? P.act_off[l] : P.col_k) * VT;
            ffj_matvec<T, S, false>(sw + P.w_off[l], sw + P.v_off[l], a.width[l], P.outp[l], in_col, out_col, nullptr, hidden,
                                    NTH, VT);
            in_col = out_col;
          }
          const T wgt = (st == 1 || st == 2) ? T(2) : T(1);
#pragma unroll
          for (int q = 0; q < S; ++q) {
            T* c = cols + q * NTH;
            for (int j = 0; j < d; ++j) c[(P.col_acc + j) * VT] += wgt * c[(P.col_k + j) * VT];
          }
          // exact trace of df/dy, one input direction at a time
          if (want_ld) {
            T div[S];
#pragma unroll
            for (int q = 0; q < S; ++q) div[q] = T(0);
            for (int j = 0; j < d; ++j) {
              T* ta = cols + P.col_ta * VT;
              T* tb = cols + P.col_tb * VT;
              if (nl == 1) {
                const int outp = P.outp[0];
                const T v = (sw + P.v_off[0])[4 * outp + j] * (sw + P.w_off[0])[j * outp + j];
#pragma unroll
                for (int q = 0; q < S; ++q) div[q] += v;
                continue;
              }
              {  // first layer: column j of W_0
                const int outp = P.outp[0];
                const T* Wt = sw + P.w_off[0] + j * outp;
                const T* gate = sw + P.v_off[0] + 4 * outp;
                const T* hc = cols + P.act_off[0] * VT;
                for (int o = 0; o < outp; ++o) {
                  const T gw = gate[o] * Wt[o];
#pragma unroll
                  for (int q = 0; q < S; ++q) {
                    const T hv = hc[o * VT + q * NTH];
                    ta[o * VT + q * NTH] = (T(1) - hv * hv) * gw;
                  }
                }
              }
              for (int l = 1; l < nl - 1; ++l) {
                ffj_matvec<T, S, true>(sw + P.w_off[l], sw + P.v_off[l], a.width[l], P.outp[l], ta, tb,
                                       cols + P.act_off[l] * VT, true, NTH, VT);
                T* tmp = ta; ta = tb; tb = tmp;
              }
              {  // last layer: output row j only
                const int l = nl - 1, outp = P.outp[l];
                const T* Wt = sw + P.w_off[l] + j;
                T acc[S];
#pragma unroll
                for (int q = 0; q < S; ++q) acc[q] = T(0);
                for (int k = 0; k < a.width[l]; ++k) {
                  const T w = Wt[k * outp];
#pragma unroll
                  for (int q = 0; q < S; ++q) acc[q] = fma(w, ta[k * VT + q * NTH], acc[q]);
                }
                const T gt = (sw + P.v_off[l])[4 * outp + j];
#pragma unroll
                for (int q = 0; q < S; ++q) div[q] += gt * acc[q];
              }
            }
#pragma unroll
            for (int q = 0; q < S; ++q) dacc[q] += wgt * div[q];
          }
        }
#pragma unroll
        for (int q = 0; q < S; ++q) {
          T* c = cols + q * NTH;
          for (int j = 0; j < d; ++j) c[(P.col_y + j) * VT] += h6 * c[(P.col_acc + j) * VT];
          logp[q] -= h6 * dacc[q];
        }
      }
    }
#pragma unroll
    for (int q = 0; q < S; ++q) {
      const long long s = tile * VT + q * NTH + tid;
      if (s < a.B) {
        for (int j = 0; j < d; ++j) gy[s * d + j] = cols[(P.col_y + j) * VT + q * NTH];
        if (want_ld) gld[s] = -logp[q];
      }
    }
  }
}

template <typename T, int S>
int launch_ffjord_s(const FfjordArgs& a, cudaStream_t st, int nt, int grid, size_t smem) {
  auto kern = gpflow_ffjord_kernel<T, S>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
    cudaGetLastError();
    return GPFLOW_ERR_CUDA;
  }
  kern<<<(unsigned)grid, nt, smem, st>>>(a);
  return cudaGetLastError() == cudaSuccess ? GPFLOW_OK : GPFLOW_ERR_CUDA;
}

template <typename T>
int launch_ffjord(const FfjordArgs& a, cudaStream_t st, int sm_count, size_t smem_limit) {
  const FfjordPlan P = ffjord_plan(a.width, a.n_linear, a.logdet != nullptr);
  const size_t wbytes = (size_t)P.weights_total * sizeof(T);
  if (wbytes + (size_t)32 * P.cols * sizeof(T) > smem_limit) return GPFLOW_ERR_UNSUPPORTED;
  int vt_max = (int)((smem_limit - wbytes) / ((size_t)P.cols * sizeof(T))) / 32 * 32;   // samples per CTA tile
  if (vt_max > kFfjMaxThreads) vt_max = kFfjMaxThreads;
  // spread small batches over the SMs
  long long per_sm = (a.B + sm_count - 1) / sm_count;
  int vt = (int)((per_sm + 31) / 32 * 32);
  if (vt > vt_max) vt = vt_max;
  if (vt < 32) vt = 32;
  // fp32, full tiles: two samples per thread (weight loads shared by both; see ffj_matvec_ob)
  static const int s2_allowed = [] { const char* e = getenv("GPFLOW_FFJORD_S2"); return e == nullptr || e[0] != '0'; }();
  const int vt64 = vt_max / 64 * 64;
  // (value-only launches; with the trace columns the tile shrinks 160 -> 128 samples and S = 2 measured slower)
  const bool s2 = sizeof(T) == 4 && s2_allowed && a.logdet == nullptr && vt64 >= 64 && vt >= vt64;
  if (s2) vt = vt64;
  const size_t smem = wbytes + (size_t)vt * P.cols * sizeof(T);
  long long grid = (a.B + vt - 1) / vt;
  if (grid > sm_count) grid = sm_count;
  if (s2) return launch_ffjord_s<T, sizeof(T) == 4 ? 2 : 1>(a, st, vt / 2, (int)grid, smem);
  return launch_ffjord_s<T, 1>(a, st, vt, (int)grid, smem);
}

}  // namespace gpflow
