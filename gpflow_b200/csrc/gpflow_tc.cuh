// gpflow_b200 -- "tc": the chunked tensor-core GP-flow path (sm_100a, fp32, no Euler penalization) for the wide shapes
// (C5: d = 64, nb = 63, H = [256, 256]; C3: d = 10, nb = 9, H = [50, 50, 50] zero-padded to 64).
//
// Same math and reference call sites as gpflow_generic.cuh:
//   VectorField.eval/.jacobian   src/gp_flows/vector_field.py:75-128     divergence_free_vector_func  velocity_field.py:20-79
//   RK4 compute_step / ode_solve ode_order_method.py:66-71, ode_solver.py:70-81   ResNet.solve / constraints res_net.py:199-317
//   Network.gp_flow network.py:156-189      Network.loss_gp_flow + backward  network.py:70-84,104-111
// but decomposed the other way round: instead of one warp per sample, a SLAB of S samples goes through every phase of a
// velocity evaluation together, so that each contraction is a real GEMM on the tcgen05 engine (gpflow_tc_gemm.cuh):
//   hidden layers      [S (d+1)] x H x H       rows = (tangent direction j, sample s), weights by TMA, shared by the slab
//   last layer         the contracted form of gpflow_generic.cuh (no Jacobian): F = h W^T, D_i = T_L[i] W_i^T, G_n = R_n W_n^T
//                      as grouped GEMMs with M = samples, one group per coordinate i / block n
//   weight gradients   K = samples (x tangent directions): reduced over the slab INSIDE the GEMM, one partial per CTA
// and everything that is not a contraction (tanh, suffix sums, the closed form A.2 and its adjoint A.3, RK4 bookkeeping,
// erf / erfinv wrapper) is a coalesced element-wise kernel over the slab.  Layouts: "RL" = row-major [row][unit] with
// row = j * S + s (what the GEMMs read and write), "CL" = column layout [index][s] for per-sample vectors.
// The tensor-core accumulator truncates when it adds (measured: error grows linearly with K), so no accumulation runs
// longer than 8 K-chunks (256) inside TMEM: longer contractions are split into segments that are added in fp32 (RN).
#pragma once
#include <string.h>

#include "gpflow_dispatch.h"
#include "gpflow_tc_gemm.cuh"

namespace gpflow {
namespace tc {

constexpr int TC_DMAX = 64;      // d <= 64: every last-layer group fits one 64-wide N tile
constexpr int TC_MAXL = GPFLOW_MAX_HIDDEN;
constexpr int TC_SEG = 8;        // K chunks per TMEM accumulation (see header)
constexpr int TC_L1_SPB = 8;     // samples per block of the first-layer kernels
#ifndef GPF_TC_L1_SPB_WIDE
#define GPF_TC_L1_SPB_WIDE 8
#endif
constexpr int TC_L1_SPB_WIDE = GPF_TC_L1_SPB_WIDE;   // tc_layer1_kernel for d > 16

struct TcDims {
  int d, nb, L, nout, P;
  int H[TC_MAXL + 2];    // H[0] = d + 1, H[1..L] hidden widths (caller's), H[L+1] = nout
  int Hp[TC_MAXL + 2];   // hidden widths zero-padded to a multiple of 64 (a padded unit is inert, see gpflow_generic.cuh)
  int woff[TC_MAXL + 1], boff[TC_MAXL + 1];   // offsets of W_li / b_li in the caller's flat parameters, li = 0..L
  int S;                 // samples per slab (multiple of 256)
  int HT;                // N tile of the hidden-unit GEMMs: 64 or 256
  int dpad;              // d rounded up to 32
  int rows3, rowsG, ldT, HTp, nbpad;   // paddings of the weight copies (see TcWeights)
};

inline int tc_round(int a, int b) { return (a + b - 1) / b * b; }

inline TcDims tc_make_dims(const GpFlowDesc* d, int S) {
  TcDims g;
  memset(&g, 0, sizeof(g));
  g.d = d->d; g.nb = d->nb; g.L = d->n_hidden; g.nout = d->d * d->nb;
  g.H[0] = g.Hp[0] = d->d + 1;
  int hmax = 0;
  for (int i = 0; i < g.L; ++i) {
    g.H[i + 1] = d->hidden[i];
    g.Hp[i + 1] = tc_round(d->hidden[i], 64);
    if (g.Hp[i + 1] > hmax) hmax = g.Hp[i + 1];
  }
  g.H[g.L + 1] = g.Hp[g.L + 1] = g.nout;
  int off = 0;
  for (int li = 0; li <= g.L; ++li) {
    g.woff[li] = off; off += g.H[li + 1] * g.H[li];
    g.boff[li] = off; off += g.H[li + 1];
  }
  g.P = off;
  g.S = S;
  g.HT = hmax <= 64 ? 64 : 256;
  g.dpad = tc_round(g.d, 32);
  g.rows3 = tc_round(g.nout, 64) + 64;
  g.rowsG = tc_round(g.nb * g.d, 64) + 64;
  g.ldT = tc_round(g.nout, 32) + 32;
  g.HTp = tc_round(g.Hp[g.L], g.HT);
  g.nbpad = tc_round(g.nb, 32);
  return g;
}

// shapes the path implements; everything else stays on the generic kernel
inline bool tc_shape_ok(const GpFlowDesc* d) {
  if (d->dtype != GPFLOW_F32 || (d->flags & GPFLOW_FLAG_EULER_PEN)) return false;
  if (d->d < 2 || d->d > TC_DMAX || d->n_hidden < 1) return false;
  for (int i = 0; i < d->n_hidden; ++i)
    if (d->hidden[i] < 1 || tc_round(d->hidden[i], 64) > 256) return false;
  return d->n_steps <= 4096;
}

// ---------------------------------------------------------------------------------------------------------------
// weight copies, prepared once per call
// ---------------------------------------------------------------------------------------------------------------
struct TcWeights {
  float* W1T;                     // [d + 1][Hp1]   W1T[j][k] = W_1[k][j]  (j = d: time column)
  float* bp[TC_MAXL + 1];         // [Hp_l] zero-padded biases, l = 1..L
  float* Ws[TC_MAXL + 1];         // l = 2..L: tf32 split of W_l, K-major [2][npad_l][Hp_{l-1}] (hi rows, then lo rows)
  float* WTs[TC_MAXL + 1];        // l = 2..L: tf32 split of W_l^T, [2][npadT_l][Hp_l]
  float* W3s;                     // [2][rows3][HpL]      rows p = i nb + n of W_{L+1}
  float* W3Gs;                    // [2][rowsG][HpL]      row n d + i = W_{L+1}[i nb + n]  (block-n rows contiguous)
  float* W3Ts;                    // [2][HTp][ldT]        W3Ts[k][p] = W_{L+1}[p][k]
  float* W3GTs;                   // [nb][2][HTp][dpad]   [n][k][i] = W_{L+1}[i nb + n][k] for i >= n, else 0
  float* W3JTs;                   // [d][2][HTp][nbpad]   [j][k][n] = W_{L+1}[j nb + n][k] for n <= min(j, nb-1), else 0
                                  // (TMA boxes start at 16-byte aligned inner coordinates only)
  const float* b3;                // [nout] (caller's parameters)
  int npad[TC_MAXL + 1], npadT[TC_MAXL + 1];
};

GPF_DEV float tc_w(const TcDims& g, const float* __restrict__ p, int li, int k, int m) {   // W_li[k][m] or 0 in the padding
  return (k < g.H[li + 1] && m < g.H[li]) ? p[g.woff[li] + k * g.H[li] + m] : 0.f;
}
GPF_DEV void tc_put_split(float* dst, size_t lo_off, size_t idx, float w) {
  float hi, lo;
  split_tf32(w, hi, lo);
  dst[idx] = hi;
  dst[lo_off + idx] = lo;
}
// one launch per copy kind (blockIdx.y selects the layer where it applies); all destinations were zero-filled
__global__ void tc_prep_first_kernel(TcDims g, const float* __restrict__ p, TcWeights w) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = (g.d + 1) * g.Hp[1];
  if (idx < n) {
    const int j = idx / g.Hp[1], k = idx % g.Hp[1];
    w.W1T[idx] = tc_w(g, p, 0, k, j);
  }
  for (int l = 1; l <= g.L; ++l)
    if (idx < g.Hp[l]) w.bp[l][idx] = idx < g.H[l] ? p[g.boff[l - 1] + idx] : 0.f;
}
__global__ void tc_prep_hidden_kernel(TcDims g, const float* __restrict__ p, TcWeights w) {
  const int l = blockIdx.y + 2;   // W_l: [Hp_l][Hp_{l-1}], flat layer index li = l - 1
  const int no = g.Hp[l], ni = g.Hp[l - 1];
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= no * ni) return;
  const int k = idx / ni, m = idx % ni;
  const float v = tc_w(g, p, l - 1, k, m);
  tc_put_split(w.Ws[l], (size_t)w.npad[l] * ni, (size_t)k * ni + m, v);
  tc_put_split(w.WTs[l], (size_t)w.npadT[l] * no, (size_t)m * no + k, v);
}
__global__ void tc_prep_last_kernel(TcDims g, const float* __restrict__ p, TcWeights w) {
  const int HL = g.Hp[g.L];
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= g.nout * HL) return;
  const int q = idx / HL, k = idx % HL;   // output row q = i nb + n
  const int i = q / g.nb, n = q % g.nb;
  const float v = tc_w(g, p, g.L, q, k);
  tc_put_split(w.W3s, (size_t)g.rows3 * HL, (size_t)q * HL + k, v);
  tc_put_split(w.W3Gs, (size_t)g.rowsG * HL, (size_t)(n * g.d + i) * HL + k, v);
  tc_put_split(w.W3Ts, (size_t)g.HTp * g.ldT, (size_t)k * g.ldT + q, v);
  if (i >= n) {
    tc_put_split(w.W3GTs + (size_t)n * 2 * g.HTp * g.dpad, (size_t)g.HTp * g.dpad, (size_t)k * g.dpad + i, v);
    tc_put_split(w.W3JTs + (size_t)i * 2 * g.HTp * g.nbpad, (size_t)g.HTp * g.nbpad, (size_t)k * g.nbpad + n, v);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// element-wise kernels of one velocity evaluation (forward)
// ---------------------------------------------------------------------------------------------------------------
// first layer (vector_field.py:84-88,109-118 for l = 1): a = W1x x + w1t t + b1, h = tanh a, sp = 1 - h^2, tangent rows
// T1[j] = sp (.) W1x[:, j] (U_1 = W1x is the same for every sample), s_j = x_j^2 - (1 - eps).  Block = 8 samples x Hp1 units.
// DB: compile-time bound on d (the per-direction loops are unrolled and predicated: at d = 10 a bound of 64 made the kernel
// instruction-bound, ncu: issue slots 53 % busy at 24 % occupancy); SPB: samples per block (fewer = more blocks in flight).
template <int DB, int SPB>
__global__ void tc_layer1_kernel(TcDims g, const float* __restrict__ W1T, const float* __restrict__ b1,
                                 const float* __restrict__ x /*CL [d][S]*/, float t, float bnd, float* __restrict__ T1 /*[(d+1) S][Hp1]*/,
                                 float* __restrict__ sp1 /*[S][Hp1]*/, float* __restrict__ sj /*CL [d][S]*/,
                                 int write_tangents /* 0: the consumers regenerate T1[j] = sp1 (.) W1x[:, j] on the fly */) {
  __shared__ float xs[DB][SPB];
  const int H = g.Hp[1], d = g.d, S = g.S;
  const int s0 = blockIdx.x * SPB, k = threadIdx.x;
  for (int e = threadIdx.x; e < d * SPB; e += blockDim.x) {
    const int j = e / SPB, q = e % SPB;
    const float xv = x[(size_t)j * S + s0 + q];
    xs[j][q] = xv;
    sj[(size_t)j * S + s0 + q] = xv * xv - bnd;
  }
  __syncthreads();
  float w[DB];
#pragma unroll
  for (int j = 0; j < DB; ++j) w[j] = j < d ? W1T[j * H + k] : 0.f;
  const float a0 = fmaf(W1T[d * H + k], t, b1[k]);
#pragma unroll
  for (int q = 0; q < SPB; ++q) {
    const int s = s0 + q;
    float a = a0;
#pragma unroll
    for (int j = 0; j < DB; ++j)
      if (j < d) a = fmaf(w[j], xs[j][q], a);
    const float h = act_tanh(a);
    const float sp = 1.f - h * h;
    sp1[(size_t)s * H + k] = sp;
    T1[((size_t)d * S + s) * H + k] = h;
    if (write_tangents) {
#pragma unroll
      for (int j = 0; j < DB; ++j)
        if (j < d) T1[((size_t)j * S + s) * H + k] = sp * w[j];
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// 64-wide nets (every Hp_l = 64, d <= DB; C3): the VALUE rows of all layers in one launch, on the CUDA cores in plain fp32.
//   h_1 = tanh(W1x x + w1t t + b1),  h_l = tanh(W_l h_{l-1} + b_l),  sp_l = 1 - h_l^2,  s_j,  [T_1[j] = sp_1 (.) W1x[:, j]]
// As GEMMs these are S x 64 x 64: 128 tiles, one per CTA, 13 us of launch / pipeline-fill latency each and two of them per
// evaluation on the critical path (the tangent GEMM of layer l needs sp_l).  Here a block takes 64 samples through the chain:
// weights transposed in shared memory ([k][unit], 16 KB per layer), activations ping-pong as [unit][sample]; thread =
// (4 units, 4 samples), per k one 16-byte load of each and 16 FMAs.
// ---------------------------------------------------------------------------------------------------------------
struct TcChainPtrs {
  float* T[TC_MAXL + 1];          // T_l: h_l goes to row block d
  float* sp[TC_MAXL + 1];
  const float* bp[TC_MAXL + 1];   // zero-padded biases
};
constexpr int TC_CH_S = 64, TC_CH_HS = 68;   // samples per block, row stride of the activation buffers
inline size_t tc_chain_smem(int L) { return ((size_t)(L - 1) * 64 * 64 + 2 * 64 * TC_CH_HS) * sizeof(float); }
template <int DB>
__global__ void __launch_bounds__(256) tc_value_chain_kernel(TcDims g, const float* __restrict__ params, const float* __restrict__ W1T,
                                                              TcChainPtrs q, const float* __restrict__ x /*CL [d][S]*/, float t, float bnd,
                                                              float* __restrict__ sj, int write_tangents) {
  extern __shared__ __align__(16) float chain_smem[];
  __shared__ float xs[DB][TC_CH_S];
  constexpr int H = 64;
  const int d = g.d, S = g.S, L = g.L, tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int s0 = blockIdx.x * TC_CH_S;
  float* Wt = chain_smem;                               // [L - 1][k][unit]
  float* hT = chain_smem + (size_t)(L - 1) * H * H;     // [2][unit][TC_CH_HS]
  for (int e = tid; e < d * TC_CH_S; e += 256) {
    const int j = e / TC_CH_S, c = e % TC_CH_S;
    const float xv = x[(size_t)j * S + s0 + c];
    xs[j][c] = xv;
    sj[(size_t)j * S + s0 + c] = xv * xv - bnd;
  }
  for (int l = 2; l <= L; ++l)
    for (int e = tid; e < H * H; e += 256) {              // e = unit * 64 + k: coalesced reads of W_l[unit][:]
      const int u = e >> 6, k = e & 63;
      Wt[(size_t)(l - 2) * H * H + k * H + u] = tc_w(g, params, l - 1, u, k);
    }
  __syncthreads();
  float a[4][4];   // [sample][unit]
  {
    float4 w1[DB];
#pragma unroll
    for (int j = 0; j < DB; ++j) w1[j] = j < d ? __ldg(reinterpret_cast<const float4*>(W1T + j * H) + tx) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float4 wt = __ldg(reinterpret_cast<const float4*>(W1T + d * H) + tx), b1 = __ldg(reinterpret_cast<const float4*>(q.bp[1]) + tx);
    const float a0[4] = {fmaf(wt.x, t, b1.x), fmaf(wt.y, t, b1.y), fmaf(wt.z, t, b1.z), fmaf(wt.w, t, b1.w)};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      a[i][0] = a0[0]; a[i][1] = a0[1]; a[i][2] = a0[2]; a[i][3] = a0[3];
    }
#pragma unroll
    for (int j = 0; j < DB; ++j)
      if (j < d) {
        const float4 xv = *reinterpret_cast<const float4*>(&xs[j][4 * ty]);
        const float xi[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          a[i][0] = fmaf(w1[j].x, xi[i], a[i][0]); a[i][1] = fmaf(w1[j].y, xi[i], a[i][1]);
          a[i][2] = fmaf(w1[j].z, xi[i], a[i][2]); a[i][3] = fmaf(w1[j].w, xi[i], a[i][3]);
        }
      }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int s = s0 + 4 * ty + i;
      float4 h, sp;
      h.x = act_tanh(a[i][0]); h.y = act_tanh(a[i][1]); h.z = act_tanh(a[i][2]); h.w = act_tanh(a[i][3]);
      sp.x = 1.f - h.x * h.x; sp.y = 1.f - h.y * h.y; sp.z = 1.f - h.z * h.z; sp.w = 1.f - h.w * h.w;
      a[i][0] = h.x; a[i][1] = h.y; a[i][2] = h.z; a[i][3] = h.w;
      reinterpret_cast<float4*>(q.sp[1] + (size_t)s * H)[tx] = sp;
      reinterpret_cast<float4*>(q.T[1] + ((size_t)d * S + s) * H)[tx] = h;
      if (write_tangents) {
#pragma unroll
        for (int j = 0; j < DB; ++j)
          if (j < d)
            reinterpret_cast<float4*>(q.T[1] + ((size_t)j * S + s) * H)[tx] =
                make_float4(sp.x * w1[j].x, sp.y * w1[j].y, sp.z * w1[j].z, sp.w * w1[j].w);
      }
    }
  }
  int cur = 0;
  for (int l = 2; l <= L; ++l, cur ^= 1) {
    float* hb = hT + (size_t)cur * H * TC_CH_HS;
#pragma unroll
    for (int u = 0; u < 4; ++u)
      *reinterpret_cast<float4*>(hb + (size_t)(4 * tx + u) * TC_CH_HS + 4 * ty) = make_float4(a[0][u], a[1][u], a[2][u], a[3][u]);
    __syncthreads();   // (the other buffer was last read two layers ago)
    const float4 bl = __ldg(reinterpret_cast<const float4*>(q.bp[l]) + tx);
#pragma unroll
    for (int i = 0; i < 4; ++i) { a[i][0] = bl.x; a[i][1] = bl.y; a[i][2] = bl.z; a[i][3] = bl.w; }
    const float* wl = Wt + (size_t)(l - 2) * H * H + 4 * tx;
    const float* hr = hb + 4 * ty;
#pragma unroll 8
    for (int k = 0; k < H; ++k) {
      const float4 w4 = *reinterpret_cast<const float4*>(wl + k * H);
      const float4 h4 = *reinterpret_cast<const float4*>(hr + k * TC_CH_HS);
      const float hi[4] = {h4.x, h4.y, h4.z, h4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        a[i][0] = fmaf(w4.x, hi[i], a[i][0]); a[i][1] = fmaf(w4.y, hi[i], a[i][1]);
        a[i][2] = fmaf(w4.z, hi[i], a[i][2]); a[i][3] = fmaf(w4.w, hi[i], a[i][3]);
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int s = s0 + 4 * ty + i;
      float4 h, sp;
      h.x = act_tanh(a[i][0]); h.y = act_tanh(a[i][1]); h.z = act_tanh(a[i][2]); h.w = act_tanh(a[i][3]);
      sp.x = 1.f - h.x * h.x; sp.y = 1.f - h.y * h.y; sp.z = 1.f - h.z * h.z; sp.w = 1.f - h.w * h.w;
      a[i][0] = h.x; a[i][1] = h.y; a[i][2] = h.z; a[i][3] = h.w;
      reinterpret_cast<float4*>(q.sp[l] + (size_t)s * H)[tx] = sp;
      reinterpret_cast<float4*>(q.T[l] + ((size_t)d * S + s) * H)[tx] = h;
    }
  }
}

// suffix sums over the tangent directions: R[n][s][k] = sum_{j >= n} s_j[s] T_L[j][s][k], n < nb.  Block = one sample.
__global__ void tc_rsum_kernel(TcDims g, const float* __restrict__ TL, const float* __restrict__ sj, float* __restrict__ R) {
  const int H = g.Hp[g.L], S = g.S, s = blockIdx.x, k = threadIdx.x;
  float acc = 0.f;
  for (int j = g.d - 1; j >= 0; --j) {
    acc = fmaf(sj[(size_t)j * S + s], TL[((size_t)j * S + s) * H + k], acc);
    if (j < g.nb) R[((size_t)j * S + s) * H + k] = acc;
  }
}

// A.2 in contracted form, part 1: per block n the sums over j >= n.  Thread = (n, s).
__global__ void tc_vel1_kernel(TcDims g, const float* __restrict__ x, const float* __restrict__ sj, const float* __restrict__ F,
                               const float* __restrict__ D, float* __restrict__ SX, float* __restrict__ FX, float* __restrict__ SD) {
  const int S = g.S;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= g.nb * S) return;
  const int n = idx / S, s = idx % S;
  // the loads of several j are independent: 4 partial sums keep them in flight (the sums are short: <= d terms)
  float sx[4] = {0.f, 0.f, 0.f, 0.f}, fx[4] = {0.f, 0.f, 0.f, 0.f}, sd[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
  for (int j = n; j < g.d; ++j) {
    const int a = (j - n) & 3;
    const float xj = x[(size_t)j * S + s];
    const size_t p = (size_t)(j * g.nb + n) * S + s;
    sx[a] += xj;
    fx[a] = fmaf(F[p], xj, fx[a]);
    sd[a] = fmaf(sj[(size_t)j * S + s], D[p], sd[a]);
  }
  SX[idx] = (sx[0] + sx[1]) + (sx[2] + sx[3]);
  FX[idx] = (fx[0] + fx[1]) + (fx[2] + fx[3]);
  SD[idx] = (sd[0] + sd[1]) + (sd[2] + sd[3]);
}
// part 2: Q_i = sum_{n <= min(i, nb-1)} [2 f_(i,n) SX_n - 2 FX_n + g_(i,n) - SD_n], v_i = s_i Q_i, followed by the RK4
// bookkeeping of the stage (ode_order_method.py:66-71).  Thread = (i, s).
//   stage sg = 0..3: ksum (+)= wk v;  sg < 3: next stage point = xcur + ak v;  sg == 3: x_next = xcur + dt/6 ksum
//   sg = -1: no bookkeeping (Lagrangian / velocity evaluations)
__global__ void tc_vel2_kernel(TcDims g, const float* __restrict__ sj, const float* __restrict__ F, const float* __restrict__ G,
                               const float* __restrict__ SX, const float* __restrict__ FX, const float* __restrict__ SD,
                               float* __restrict__ Q, float* __restrict__ v, int sg, float dt, const float* __restrict__ xcur,
                               float* __restrict__ ksum, float* __restrict__ ynext) {
  const int S = g.S;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= g.d * S) return;
  const int i = idx / S, s = idx % S;
  const int nmax = i < g.nb - 1 ? i : g.nb - 1;
  float qa[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
  for (int n = 0; n <= nmax; ++n) {
    const size_t p = (size_t)(i * g.nb + n) * S + s;
    const size_t ns = (size_t)n * S + s;
    qa[n & 3] += 2.f * F[p] * SX[ns] - 2.f * FX[ns] + G[p] - SD[ns];
  }
  const float q = (qa[0] + qa[1]) + (qa[2] + qa[3]);
  const float vi = sj[idx] * q;
  Q[idx] = q;
  v[idx] = vi;
  if (sg < 0) return;
  const float wk = (sg == 1 || sg == 2) ? 2.f : 1.f;
  const float ks = sg == 0 ? vi : fmaf(wk, vi, ksum[idx]);
  ksum[idx] = ks;
  if (sg < 3) ynext[idx] = fmaf(sg == 2 ? dt : dt * 0.5f, vi, xcur[idx]);
  else ynext[idx] = fmaf(dt / 6.f, ks, xcur[idx]);
}

// ---------------------------------------------------------------------------------------------------------------
// element-wise kernels of the reverse sweep of one velocity evaluation (formulas: header of Generic<T>)
// ---------------------------------------------------------------------------------------------------------------
// cotangent of the stage velocity (A.5): vbar = wk xb + ak ybprev; Qb = vbar s.  Thread = (i, s).
__global__ void tc_vbar_kernel(TcDims g, float wk, float ak, const float* __restrict__ xb, const float* __restrict__ ybprev,
                               const float* __restrict__ sj, float* __restrict__ vbar, float* __restrict__ Qb) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= g.d * g.S) return;
  const float vb = wk * xb[idx] + (ak != 0.f ? ak * ybprev[idx] : 0.f);
  vbar[idx] = vb;
  Qb[idx] = vb * sj[idx];
}
// SQ_n = sum_{i >= n} Qb_i, SXb_n = sum_{i >= n} 2 f_(i,n) Qb_i.  Thread = (n, s).
__global__ void tc_vb1_kernel(TcDims g, const float* __restrict__ Qb, const float* __restrict__ F, float* __restrict__ SQ,
                              float* __restrict__ SXb) {
  const int S = g.S;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= g.nb * S) return;
  const int n = idx / S, s = idx % S;
  float sq[4] = {0.f, 0.f, 0.f, 0.f}, sxb[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
  for (int i = n; i < g.d; ++i) {
    const int a = (i - n) & 3;
    const float qb = Qb[(size_t)i * S + s];
    sq[a] += qb;
    sxb[a] = fmaf(2.f * F[(size_t)(i * g.nb + n) * S + s], qb, sxb[a]);
  }
  SQ[idx] = (sq[0] + sq[1]) + (sq[2] + sq[3]);
  SXb[idx] = (sxb[0] + sxb[1]) + (sxb[2] + sxb[3]);
}
// xbar_j (first part) = sum_{n <= min(j, nb-1)} [SXb_n - 2 SQ_n f_(j,n)] + 2 x_j (vbar_j Q_j - sum_n SQ_n D_(j,n));
// the CR . T_L[j] term of sb_j and the first-layer term are added by tc_tb_last_kernel / tc_layer1_bwd_kernel.
__global__ void tc_vb2_kernel(TcDims g, const float* __restrict__ x, const float* __restrict__ vbar, const float* __restrict__ Q,
                              const float* __restrict__ F, const float* __restrict__ D, const float* __restrict__ SQ,
                              const float* __restrict__ SXb, float* __restrict__ ybar) {
  const int S = g.S;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= g.d * S) return;
  const int j = idx / S, s = idx % S;
  const int nmax = j < g.nb - 1 ? j : g.nb - 1;
  float xb[4] = {0.f, 0.f, 0.f, 0.f}, sd[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
  for (int n = 0; n <= nmax; ++n) {
    const size_t p = (size_t)(j * g.nb + n) * S + s;
    const float sq = SQ[(size_t)n * S + s];
    xb[n & 3] += SXb[(size_t)n * S + s] - 2.f * sq * F[p];
    sd[n & 3] = fmaf(sq, D[p], sd[n & 3]);
  }
  const float sb = vbar[idx] * Q[idx] - ((sd[0] + sd[1]) + (sd[2] + sd[3]));
  ybar[idx] = fmaf(2.f * x[idx], sb, (xb[0] + xb[1]) + (xb[2] + xb[3]));
}
// last hidden layer, reverse: Tb_L[j] = s_j CR_{min(j, nb-1)} + TbD[j] (TbD already carries -s_j sum_n SQ_n W[(j,n)]),
// CR_n = sum_{n' <= n} Rb_n';  sb_j += CR . T_L[j];  Ub_L[j] = sp (.) Tb_L[j];
// ab_L = hb sp + (sum_j Tb_L[j] (.) T_L[j]) (-2 h)   [sigma'' = -2 h sigma', T = sigma' U];  hb = sum of the K segments.
// Block = one sample x HpL units.  Ub (in: TbD, out: Ub_L rows, ab_L in row block d).
__global__ void tc_tb_last_kernel(TcDims g, const float* __restrict__ TL /*[(d+1) S][H], block d = h_L*/, const float* __restrict__ spL,
                                  const float* __restrict__ Rb, const float* __restrict__ hbp, int nseg_hb,
                                  const float* __restrict__ sj, const float* __restrict__ x, float* __restrict__ Ub,
                                  float* __restrict__ ybar, int raw /* L == 1: leave Tb_1 / hb_1 for tc_layer1_bwd_kernel */) {
  __shared__ float sbd[8][TC_DMAX];   // per-warp partial dots (HpL <= 256: at most 8 warps), summed in a fixed order
  const int H = g.Hp[g.L], S = g.S, d = g.d, s = blockIdx.x, k = threadIdx.x;
  const int lane = k & 31, warp = k >> 5, nwarp = H >> 5;
  const float sp = spL[(size_t)s * H + k];
  float cr = 0.f, acc = 0.f;
  for (int j = 0; j < d; ++j) {
    if (j < g.nb) cr += Rb[((size_t)j * S + s) * H + k];
    const size_t o = ((size_t)j * S + s) * H + k;
    const float tl = TL[o];
    const float tb = fmaf(sj[(size_t)j * S + s], cr, Ub[o]);
    Ub[o] = raw ? tb : sp * tb;
    acc = fmaf(tb, tl, acc);
    const float dot = warp_sum(cr * tl);
    if (lane == 0) sbd[warp][j] = dot;
  }
  float hb = 0.f;
  for (int q = 0; q < nseg_hb; ++q) hb += hbp[((size_t)q * S + s) * H + k];
  const float h = TL[((size_t)d * S + s) * H + k];
  Ub[((size_t)d * S + s) * H + k] = raw ? hb : fmaf(hb, sp, acc * (-2.f * h));
  __syncthreads();
  if (k < d) {
    float t = 0.f;
    for (int q = 0; q < nwarp; ++q) t += sbd[q][k];
    ybar[(size_t)k * S + s] += 2.f * x[(size_t)k * S + s] * t;
  }
}
// inner hidden layer 2 <= l < L, reverse (A.4): in Ub: Tb_l rows and hb_l (row block d); out: Ub_l = sp (.) Tb_l,
// ab_l = hb sp + (sum_j Tb_l[j] (.) T_l[j]) (-2 h_l)
__global__ void tc_tb_mid_kernel(TcDims g, int l, const float* __restrict__ Tl, const float* __restrict__ spl, float* __restrict__ Ub) {
  const int H = g.Hp[l], S = g.S, d = g.d, s = blockIdx.x, k = threadIdx.x;
  const float sp = spl[(size_t)s * H + k];
  float acc = 0.f;
  for (int j = 0; j < d; ++j) {
    const size_t o = ((size_t)j * S + s) * H + k;
    const float tb = Ub[o];
    Ub[o] = sp * tb;
    acc = fmaf(tb, Tl[o], acc);
  }
  const size_t o = ((size_t)d * S + s) * H + k;
  Ub[o] = fmaf(Ub[o], sp, acc * (-2.f * Tl[o]));
}
// first layer, reverse (A.4, l = 1), part a.  In Ub: Tb_1 rows and hb_1 (row block d).
//   ab_1 = hb sp + (sum_j Tb_1[j] W1x[:, j]) (-2 h sp)   -> written over hb_1 (row block d of Ub)
//   xbar_j += W1x[:, j] . ab_1
// Block = one sample x Hp1 units (W1x is read through L1: 64 KB shared by every block).
constexpr int TC_L1B_SPB = 1;
__global__ void tc_layer1_bwd_a_kernel(TcDims g, const float* __restrict__ W1T, const float* __restrict__ T1, const float* __restrict__ sp1,
                                       float* __restrict__ Ub, float* __restrict__ ybar) {
  __shared__ float xbs[8][TC_DMAX];   // per-warp partial dots, summed in a fixed order
  const int H = g.Hp[1], d = g.d, S = g.S;
  const int k = threadIdx.x, lane = k & 31, warp = k >> 5, nwarp = H >> 5;
  const int s = blockIdx.x;
  const float sp = sp1[(size_t)s * H + k];
  const float h = T1[((size_t)d * S + s) * H + k];
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
  for (int j = 0; j < d; ++j) acc[j & 3] = fmaf(Ub[((size_t)j * S + s) * H + k], __ldg(W1T + j * H + k), acc[j & 3]);
  const size_t o = ((size_t)d * S + s) * H + k;
  const float ab = fmaf(Ub[o], sp, ((acc[0] + acc[1]) + (acc[2] + acc[3])) * (-2.f * h * sp));
  Ub[o] = ab;
  for (int j = 0; j < d; ++j) {
    const float dot = warp_sum(__ldg(W1T + j * H + k) * ab);
    if (lane == 0) xbs[warp][j] = dot;
  }
  __syncthreads();
  if (k < d) {
    float tt = 0.f;
    for (int r = 0; r < nwarp; ++r) tt += xbs[r][k];
    ybar[(size_t)k * S + s] += tt;
  }
}
// part b: dW1x[:, j] += sum_s (ab_1 x_j + sp Tb_1[j]),  dw1t += sum_s ab_1 t,  db1 += sum_s ab_1.
// Grid (Hp1 / 32, ceil(d / 8), nz): a block reduces 8 tangent directions over its sample range (8 warps stride the
// samples) and ADDS the result to its entry of the per-split planes [nz][Hp1][d + 2] (one owner thread per entry, calls
// are stream-ordered: deterministic); the planes live for the whole call (zeroed with the gradient accumulators) and are
// summed once, in a fixed order, by tc_reduce_add_kernel -- not once per evaluation (240 single-block launches per C3 step).
__global__ void __launch_bounds__(256) tc_layer1_bwd_b_kernel(TcDims g, const float* __restrict__ sp1, const float* __restrict__ Ub,
                                                               const float* __restrict__ x, float t, int s_per_z, float* __restrict__ part) {
  __shared__ float red[8][10][33];
  const int H = g.Hp[1], d = g.d, S = g.S;
  const int lane = threadIdx.x, wy = threadIdx.y, k = blockIdx.x * 32 + lane, j0 = blockIdx.y * 8;
  const int s_lo = blockIdx.z * s_per_z, s_hi = s_lo + s_per_z < S ? s_lo + s_per_z : S;
  float acc[8], gt = 0.f, gb = 0.f;
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
#pragma unroll 4
  for (int s = s_lo + wy; s < s_hi; s += 8) {
    const float sp = sp1[(size_t)s * H + k];
    const float ab = Ub[((size_t)d * S + s) * H + k];
    gt = fmaf(ab, t, gt);
    gb += ab;
#pragma unroll
    for (int e = 0; e < 8; ++e)
      if (j0 + e < d) acc[e] += fmaf(ab, x[(size_t)(j0 + e) * S + s], sp * Ub[((size_t)(j0 + e) * S + s) * H + k]);
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) red[wy][e][lane] = acc[e];
  red[wy][8][lane] = gt;
  red[wy][9][lane] = gb;
  __syncthreads();
  if (wy == 0) {
    float* o = part + ((size_t)blockIdx.z * H + k) * (d + 2);
    for (int e = 0; e < 10; ++e) {
      float v = 0.f;
      for (int r = 0; r < 8; ++r) v += red[r][e][lane];
      if (e < 8) { if (j0 + e < d) o[j0 + e] += v; }
      else if (blockIdx.y == 0) o[d + (e - 8)] += v;
    }
  }
}
// The same with float4 loads, for first layers whose padded width is a multiple of 64.  The kernel above keeps few warps on a
// serial chain of load batches (ncu at C3: 67 us per launch, long_scoreboard 84 % of the stall samples, 22 % occupancy).
// Here a block owns TC_L1B_CH = 32 samples, 64 units (blockIdx.z) and DB = 16 directions (blockIdx.y) of one plane:
// thread = (4 units, 1 of 16 sample rows), all the loads of a thread (2 samples x (DB + 2)) are independent; the 16 sample
// rows are summed by one shuffle and a fixed-order pass over the 8 warps.  Grid (S / 32, ceil(d / 16), Hp1 / 64).
constexpr int TC_L1B_CH = 32;
template <int DB>
__global__ void __launch_bounds__(256) tc_layer1_bwd_b_small_kernel(TcDims g, const float* __restrict__ sp1, const float* __restrict__ Ub,
                                                                     const float* __restrict__ x, float t, float* __restrict__ part) {
  __shared__ float xs[DB][TC_L1B_CH];
  __shared__ float4 red[8][DB + 2][16];
  const int H = g.Hp[1], d = g.d, S = g.S, tid = threadIdx.x, k4 = tid & 15, sr = tid >> 4, lane = tid & 31, warp = tid >> 5;
  const int s0 = blockIdx.x * TC_L1B_CH, j0 = blockIdx.y * DB, kb = blockIdx.z * 64;
  const int dj = d - j0 < DB ? d - j0 : DB;   // directions of this block
  for (int e = tid; e < dj * TC_L1B_CH; e += 256) {
    const int j = e / TC_L1B_CH, q = e % TC_L1B_CH;
    xs[j][q] = x[(size_t)(j0 + j) * S + s0 + q];
  }
  __syncthreads();
  float4 acc[DB], gb = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int j = 0; j < DB; ++j) acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int i = 0; i < TC_L1B_CH / 16; ++i) {
    const int q = sr + 16 * i, s = s0 + q;
    const float4 sp = __ldg(reinterpret_cast<const float4*>(sp1 + (size_t)s * H + kb) + k4);
    const float4 ab = __ldg(reinterpret_cast<const float4*>(Ub + ((size_t)d * S + s) * H + kb) + k4);
    gb.x += ab.x; gb.y += ab.y; gb.z += ab.z; gb.w += ab.w;
#pragma unroll
    for (int j = 0; j < DB; ++j)
      if (j < dj) {
        const float4 u = __ldg(reinterpret_cast<const float4*>(Ub + ((size_t)(j0 + j) * S + s) * H + kb) + k4);
        const float xj = xs[j][q];
        acc[j].x += fmaf(ab.x, xj, sp.x * u.x); acc[j].y += fmaf(ab.y, xj, sp.y * u.y);
        acc[j].z += fmaf(ab.z, xj, sp.z * u.z); acc[j].w += fmaf(ab.w, xj, sp.w * u.w);
      }
  }
  auto fold = [&](float4 v, int o) {   // the two sample rows of a warp, then its slot
    v.x += __shfl_xor_sync(0xffffffffu, v.x, 16); v.y += __shfl_xor_sync(0xffffffffu, v.y, 16);
    v.z += __shfl_xor_sync(0xffffffffu, v.z, 16); v.w += __shfl_xor_sync(0xffffffffu, v.w, 16);
    if (lane < 16) red[warp][o][k4] = v;
  };
#pragma unroll
  for (int j = 0; j < DB; ++j)
    if (j < dj) fold(acc[j], j);
  fold(make_float4(gb.x * t, gb.y * t, gb.z * t, gb.w * t), DB);
  fold(gb, DB + 1);
  __syncthreads();
  const float* rf = reinterpret_cast<const float*>(&red[0][0][0]);
  const int nout = dj + (blockIdx.y == 0 ? 2 : 0);   // the time-column and bias sums are written by the first direction tile
  for (int e = tid; e < nout * 64; e += 256) {
    const int oo = e / 64, k = e % 64;
    const int slot = oo < dj ? oo : DB + (oo - dj), col = oo < dj ? j0 + oo : d + (oo - dj);
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) v += rf[(w * (DB + 2) + slot) * 64 + k];
    part[((size_t)blockIdx.x * H + kb + k) * (d + 2) + col] += v;
  }
}
// acc[e] += sum_z part[z][e]  (fixed order)
__global__ void tc_reduce_add_kernel(int nz, int n, const float* __restrict__ part, float* __restrict__ acc) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  float s = 0.f;
  for (int z = 0; z < nz; ++z) s += part[(size_t)z * n + e];
  acc[e] += s;
}
// bias gradients of the hidden layers: partial column sums part[z][k] = sum_{rows of split z} ab[r][k].
// Grid (H / 32, nz), block (32, 8).
__global__ void tc_colsum_kernel(int S, int H, int s_per_z, const float* __restrict__ ab, float* __restrict__ part) {
  __shared__ float red[8][33];
  const int k = blockIdx.x * 32 + threadIdx.x, y = threadIdx.y;
  const int s_lo = blockIdx.y * s_per_z, s_hi = s_lo + s_per_z < S ? s_lo + s_per_z : S;
  float s = 0.f;
  if (k < H)
    for (int r = s_lo + y; r < s_hi; r += 8) s += ab[(size_t)r * H + k];
  red[y][threadIdx.x] = s;
  __syncthreads();
  if (y == 0 && k < H) {
    float t = 0.f;
    for (int q = 0; q < 8; ++q) t += red[q][threadIdx.x];
    part[(size_t)blockIdx.y * H + k] += t;   // persistent per-split plane, see tc_layer1_bwd_b_kernel
  }
}
// db_{L+1}[(i,n)] = sum_s fb_(i,n)[s],  fb = 2 (Qb_i SX_n - x_i SQ_n), kept rows only: partial sums over 256-sample
// splits, part[z][i nb + n].  Grid (ceil(nb / 16), ceil(d / 16), S / chunk), block 16 x 16 (i, n); chunk = 64 or 256.
__global__ void tc_db3_kernel(TcDims g, const float* __restrict__ Qb, const float* __restrict__ x, const float* __restrict__ SX,
                              const float* __restrict__ SQ, int chunk, float* __restrict__ part) {
  __shared__ float a[4][16][65];
  const int S = g.S, i = blockIdx.y * 16 + threadIdx.y, n = blockIdx.x * 16 + threadIdx.x;
  const int tid = threadIdx.y * 16 + threadIdx.x;
  float acc = 0.f;
  for (int s0 = blockIdx.z * chunk; s0 < blockIdx.z * chunk + chunk; s0 += 64) {
    for (int e = tid; e < 4 * 16 * 64; e += 256) {
      const int arr = e / 1024, r = (e / 64) % 16, c = e % 64;
      const int row = (arr < 2 ? blockIdx.y : blockIdx.x) * 16 + r;
      const int lim = arr < 2 ? g.d : g.nb;
      const float* src = arr == 0 ? Qb : arr == 1 ? x : arr == 2 ? SX : SQ;
      a[arr][r][c] = row < lim ? src[(size_t)row * S + s0 + c] : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int c = 0; c < 64; ++c)
      acc += a[0][threadIdx.y][c] * a[2][threadIdx.x][c] - a[1][threadIdx.y][c] * a[3][threadIdx.x][c];
    __syncthreads();
  }
  if (i < g.d && n < g.nb && n <= i) part[(size_t)blockIdx.z * g.nout + i * g.nb + n] += 2.f * acc;   // persistent plane
}

// ---------------------------------------------------------------------------------------------------------------
// wrapper, time stepping and reductions over the slab
// ---------------------------------------------------------------------------------------------------------------
// g0 = erf(x / (sqrt2 sigma)) (network.py:174) or the raw cube-space point; samples beyond the batch are 0
__global__ void tc_prologue_kernel(TcDims g, const float* __restrict__ x_in, long long b0, int nvalid, int wrap, float sq2s,
                                   float* __restrict__ x0) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= g.d * g.S) return;
  const int s = idx / g.d, j = idx % g.d;
  float v = 0.f;
  if (s < nvalid) {
    v = x_in[(b0 + s) * g.d + j];
    if (wrap) v = erf_real(v / sq2s);
  }
  x0[(size_t)j * g.S + s] = v;
}
// y = sqrt2 sigma erfinv(clip(g_T)) (network.py:178-182), log|det| (:183-185), OT term and adjoint seed (A.6)
// mode: 0 forward op, 1 train step (seed -2 (x - y) loss_scale), 2 backward op (seed = grad_y)
__global__ void tc_epilogue_kernel(TcDims g, const float* __restrict__ gT, const float* __restrict__ x_in,
                                   const float* __restrict__ x_data, const float* __restrict__ grad_y, long long b0, int nvalid,
                                   int wrap, float sigma, float loss_scale, int mode, float* __restrict__ y_out,
                                   float* __restrict__ logdet_out, float* __restrict__ xb, float* __restrict__ ot_s) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= g.S) return;
  const int d = g.d, S = g.S;
  const float sq2s = 1.4142135623730951f * sigma;
  float ny = 0.f, nx = 0.f, ot = 0.f;
  for (int j = 0; j < d; ++j) {
    float seed = 0.f;
    if (s < nvalid) {
      const float gv = gT[(size_t)j * S + s];
      float yv = gv;
      bool inband = true;
      if (wrap) {
        const float lim = 1.f - Eps<float>::value;
        const float gc = gv < -lim ? -lim : (gv > lim ? lim : gv);
        inband = (gv >= -lim) && (gv <= lim);
        yv = sq2s * erfinv_real(gc);
      }
      const size_t o = (size_t)(b0 + s) * d + j;
      const float xi = x_in[o];
      ny += yv * yv;
      nx += xi * xi;
      if (y_out != nullptr) y_out[o] = yv;
      if (mode != 0) {
        float yb;
        if (mode == 1) {
          const float diff = x_data[o] - yv;
          ot += diff * diff;
          yb = -2.f * diff * loss_scale;
        } else {
          yb = grad_y[o];
        }
        if (wrap) {
          const float dydg = 1.2533141373155001f * sigma * exp_real(yv * yv / (2.f * sigma * sigma));
          seed = inband ? yb * dydg : 0.f;
        } else {
          seed = yb;
        }
      }
    }
    if (mode != 0) xb[(size_t)j * S + s] = seed;
  }
  if (s < nvalid && logdet_out != nullptr) logdet_out[b0 + s] = (ny - nx) / (2.f * sigma * sigma);
  if (mode == 1) ot_s[s] = ot;
}
// after the reverse sweep of a stage: ybsum (+)= ybar; ybprev = ybar;  after stage 0: xb += ybsum
__global__ void tc_stage_bwd_kernel(TcDims g, int sg, const float* __restrict__ ybar, float* __restrict__ ybsum,
                                    float* __restrict__ ybprev, float* __restrict__ xb) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= g.d * g.S) return;
  const float yb = ybar[idx];
  const float sum = sg == 3 ? yb : ybsum[idx] + yb;
  ybsum[idx] = sum;
  ybprev[idx] = yb;
  if (sg == 0) xb[idx] += sum;
}
__global__ void tc_gradx_kernel(TcDims g, const float* __restrict__ xb, const float* __restrict__ x_in, long long b0, int nvalid,
                                int wrap, float sigma, float* __restrict__ grad_x_out) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= g.d * g.S) return;
  const int s = idx / g.d, j = idx % g.d;
  if (s >= nvalid) return;
  const size_t o = (size_t)(b0 + s) * g.d + j;
  float gx = xb[(size_t)j * g.S + s];
  if (wrap) {
    const float xi = x_in[o];
    gx *= 0.7978845608028654f / sigma * exp_real(-xi * xi / (2.f * sigma * sigma));
  }
  grad_x_out[o] = gx;
}
// CL [d][S] -> row-major rows of the caller's [.., B, d] arrays (trajectory states, velocity output)
__global__ void tc_export_kernel(TcDims g, const float* __restrict__ cl, long long b0, int nvalid, float* __restrict__ out) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= g.d * g.S) return;
  const int s = idx / g.d, j = idx % g.d;
  if (s < nvalid) out[(b0 + s) * g.d + j] = cl[(size_t)j * g.S + s];
}
// per-sample 1/2 |v|^2 for the Lagrangian (res_net.py:292-294,360-361), accumulated over the steps
__global__ void tc_lagr_kernel(TcDims g, const float* __restrict__ v, int nvalid, int first, float* __restrict__ lg_s) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= g.S) return;
  float a = 0.f;
  if (s < nvalid)
    for (int j = 0; j < g.d; ++j) { const float q = v[(size_t)j * g.S + s]; a = fmaf(0.5f * q, q, a); }
  lg_s[s] = first ? a : lg_s[s] + a;
}
// fixed-order sum of a per-sample vector into one accumulator (single block: deterministic)
__global__ void tc_sum_kernel(int n, const float* __restrict__ v, float* __restrict__ acc) {
  __shared__ float red[256];
  float s = 0.f;
  for (int i = threadIdx.x; i < n; i += 256) s += v[i];
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) acc[0] += red[0];
}

}  // namespace tc
}  // namespace gpflow
