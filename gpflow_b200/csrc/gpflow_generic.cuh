// gpflow_b200 -- generic GP-flow kernels (sm_100a): any d, nb, depth and widths, fp32 / fp64.
//
// One WARP per sample; every phase is "owner computes": lanes stride over the outputs of the phase (hidden
// units k, input units m, coordinates j), read the previous phase's results from the warp's scratch region
// (global memory, L1/L2 resident) and __syncwarp() separates phases.  No atomics: each warp owns one partial
// vector [ot, reg, lagr | parameter gradients] that gpflow_finalize_kernel sums in a fixed order.
// Activations use the layout [j][k] (tangent direction major) so that lanes over k or m are coalesced; the
// forward contractions read a transposed copy of the weights (lanes over output units), the reverse ones the
// original nn.Linear layout (lanes over input units).
//
// Same math, same reference call sites as gpflow_small.cuh; scalar type S is T or
// Dual<T> (Euler penalization by reverse-over-forward, see gpflow_common.cuh).  For S = float the hidden-layer
// contractions run on the tensor cores (3xTF32 mma.sync, gpflow_generic_mma.cuh); the last layer and the closed
// form A.2 are evaluated in contracted form (3 dot products per kept output row, no Jacobian), see Generic<T>.
#pragma once
#include "../../include/gpflow_b200.h"
#include "gpflow_common.cuh"
#include <string.h>
#include <stdlib.h>
#include "gpflow_small.cuh"  // MODE_*, gpflow_finalize_kernel
#include "gpflow_generic_mma.cuh"
#include <type_traits>

namespace gpflow {

constexpr int GMAXL = GPFLOW_MAX_HIDDEN + 1;  // linear layers
constexpr int kGenericMaxThreads = 512;   // register allocation bounded for 16 warps per SM (128 registers per thread)
constexpr int kGenericSmemBudget = 226 * 1024;

struct GDims {
  int d, nb, L, nout, P;
  int Ps;               // P rounded up to 4: stride of the tf32-split weight copies (keeps them 16-byte aligned)
  int n[GMAXL + 1];     // widths: n[0] = d+1, n[1..L] hidden, n[L+1] = nout
  int woff[GMAXL];      // offset of W_li in the flat parameters, li = 0..L  (row-major [n[li+1]][n[li]])
  int boff[GMAXL];
};
struct GLayout {         // offsets (elements of the scalar type) inside one warp's scratch
  int h[GMAXL + 1], sp[GMAXL + 1], U[GMAXL + 1], T[GMAXL + 1], hb[GMAXL + 1], Tb[GMAXL + 1];  // hidden li = 1..L
  int f, G, D, fb, R;    // last layer in contracted form: f_p, g_p, D_p, cotangent of f, R_n[m]
  int x, s, Q, Qb, sb, v, vbar, xbar, SX, SQ, CN, SXb;
  int total;
};

inline GDims make_gdims(const GpFlowDesc* d) {
  GDims g;
  memset(&g, 0, sizeof(g));
  g.d = d->d; g.nb = d->nb; g.L = d->n_hidden; g.nout = d->d * d->nb;
  g.n[0] = d->d + 1;
  for (int i = 0; i < g.L; ++i) g.n[i + 1] = d->hidden[i];
  g.n[g.L + 1] = g.nout;
  int off = 0;
  for (int li = 0; li <= g.L; ++li) {
    g.woff[li] = off; off += g.n[li + 1] * g.n[li];
    g.boff[li] = off; off += g.n[li + 1];
  }
  g.P = off;
  g.Ps = (off + 3) / 4 * 4;
  return g;
}
inline GLayout make_glayout(const GDims& g) {
  GLayout l;
  memset(&l, 0, sizeof(l));
  int off = 0;
  auto take = [&](int n) { int o = off; off += (n + 3) / 4 * 4; return o; };
  for (int li = 1; li <= g.L; ++li) {
    l.h[li] = take(g.n[li]); l.sp[li] = take(g.n[li]); l.U[li] = take(g.n[li] * g.d); l.T[li] = take(g.n[li] * g.d);
    l.hb[li] = take(g.n[li]); l.Tb[li] = take(g.n[li] * g.d);
  }
  l.f = take(g.nout); l.G = take(g.nout); l.D = take(g.nout); l.fb = take(g.nout); l.R = take(g.nb * g.n[g.L]);
  l.x = take(g.d); l.s = take(g.d); l.Q = take(g.d); l.Qb = take(g.d); l.sb = take(g.d); l.v = take(g.d);
  l.vbar = take(g.d); l.xbar = take(g.d); l.SX = take(g.nb); l.SQ = take(g.nb); l.CN = take(g.nb); l.SXb = take(g.nb);
  l.total = off;
  return l;
}

template <typename T>
struct GArgs {
  GDims g;
  GLayout lay;
  const T* params;       // flat [P]
  const T* params_t;     // transposed weights: Wt_li [n[li]][n[li+1]] at woff[li] (biases not copied)
  const T* w_split;      // fp32: [W_hi | W_lo | Wt_hi | Wt_lo], P floats each: tf32 hi / lo parts of the weights and of
                         // their transposes (3xTF32 operands prepared once per call); nullptr otherwise
  const T* x_in; const T* x_data; const T* grad_y; const T* reg_cotangent; const T* zw;
  T* y_out; T* logdet_out; T* traj_out; T* grad_x_out; T* v_out;
  T* partials;           // [nwarps][3 + P]
  char* scratch;         // per-warp scratch, stride scratch_stride bytes (global mode)
  size_t scratch_stride;
  int grad_red;          // 1: gradient partials accumulated with vector `red` instead of read-modify-write
  int cta_align;         // 1: named barrier between the CTA's active warps before every stage evaluation
  int on_chip;           // 1: the parameters and every warp's scratch live in the CTA's shared memory
  int vec4;              // 1: fp32, last hidden width and the last layer's parameter offset multiples of 4, 16-byte
                         //    aligned parameters: the last-layer passes use 128-bit loads (4 input units per lane)
  unsigned smem_scratch_off;  // byte offset of warp 0's scratch inside dynamic shared memory (on-chip mode)
  long long B;
  int n_steps; int npart; unsigned flags;
  T dt, t0, dt_constraint, sigma, euler_coef, loss_scale, t_eval;
};

// ---- zero-padding of the hidden widths to the aligned tensor-core shape (fp32, no Euler) ----
// A hidden unit with zero incoming weights, zero bias and zero outgoing weights is exactly inert (h = tanh(0) = 0, its
// tangent row is 0, every adjoint that reaches it is multiplied by a zero weight), so a network whose hidden widths are
// rounded up to a multiple of 64 computes the same flow and the same gradients for the original parameters.  The aligned
// mma fast paths (pre-split weights, 128-bit operand loads, no bounds predicates) and the 128-bit last-layer passes need
// exactly that shape; measured at C3 (d=10, nb=9, H=[50,50,50] -> [64,64,64], B = 2^14): 242 -> 184 ms per train step.
// Applied when it costs at most 30 % more units per layer; GPFLOW_GENERIC_PAD=0 disables it (A/B measurements).
inline bool generic_padded_desc(const GpFlowDesc* d, GpFlowDesc* out) {
  static const int allowed = [] { const char* e = getenv("GPFLOW_GENERIC_PAD"); return e == nullptr || e[0] != '0'; }();
  *out = *d;
  if (!allowed || d->dtype != GPFLOW_F32 || d->n_hidden < 2 || (d->flags & GPFLOW_FLAG_EULER_PEN) || !GPF_GENERIC_MMA)
    return false;
  bool changed = false;
  for (int l = 0; l < d->n_hidden; ++l) {
    const int h = d->hidden[l], hp = (h + 63) / 64 * 64;
    if (hp == h) continue;
    if (10 * hp > 13 * h) return false;
    out->hidden[l] = hp;
    changed = true;
  }
  return changed;
}

// scatter the caller's flat parameters into the (pre-zeroed) padded layout; map[i] = padded index of parameter i
template <typename T>
__global__ void gpflow_pad_params_kernel(GDims g, GDims gp, const T* __restrict__ p, T* __restrict__ pp, int* __restrict__ map) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.P) return;
  int idx = 0;
  for (int li = 0; li <= g.L; ++li) {
    const int nin = g.n[li], no = g.n[li + 1];
    if (i >= g.woff[li] && i < g.woff[li] + nin * no) {
      const int e = i - g.woff[li], k = e / nin, m = e - k * nin;
      idx = gp.woff[li] + k * gp.n[li] + m;
      break;
    }
    if (i >= g.boff[li] && i < g.boff[li] + no) {
      idx = gp.boff[li] + (i - g.boff[li]);
      break;
    }
  }
  pp[idx] = p[i];
  map[i] = idx;
}

// gpflow_finalize_kernel with the gradients gathered back from the padded layout
template <typename T>
__global__ void gpflow_finalize_mapped_kernel(const T* __restrict__ partials, int n_cta, int npart, int count,
                                              const int* __restrict__ map, T* __restrict__ scalars_out,
                                              T* __restrict__ grads_out) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= count) return;
  const int src = e < 3 ? e : 3 + map[e - 3];
  T s = T(0);
  for (int c = 0; c < n_cta; ++c) s += partials[(size_t)c * npart + src];
  if (e < 3) {
    if (scalars_out != nullptr) scalars_out[e] = s;
  } else if (grads_out != nullptr) {
    grads_out[e - 3] = s;
  }
}

template <typename T>
__global__ void gpflow_transpose_params_kernel(GDims g, const T* __restrict__ p, T* __restrict__ pt, T* __restrict__ split) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.P) return;
  int it = i;  // index of p[i] in the transposed copy (biases stay in place)
  for (int li = 0; li <= g.L; ++li) {
    const int nin = g.n[li], no = g.n[li + 1];
    if (i >= g.woff[li] && i < g.woff[li] + nin * no) {
      const int e = i - g.woff[li], k = e / nin, m = e % nin;
      it = g.woff[li] + m * no + k;
      break;
    }
  }
  pt[it] = p[i];
  if constexpr (std::is_same<T, float>::value) {
    if (split != nullptr) {
      uint32_t hi, lo;
      tf32_split(p[i], hi, lo);
      split[i] = __uint_as_float(hi);
      split[(size_t)g.Ps + i] = __uint_as_float(lo);
      split[2 * (size_t)g.Ps + it] = __uint_as_float(hi);
      split[3 * (size_t)g.Ps + it] = __uint_as_float(lo);
    }
  }
}

// flat parameters as the kernels read them: the CTA's shared-memory copy in on-chip mode, else the caller's buffer
template <typename T>
GPF_DEV const T* gparams(const GArgs<T>& a) {
  extern __shared__ __align__(16) unsigned char g_dyn_smem[];
  return a.on_chip ? reinterpret_cast<const T*>(g_dyn_smem) : a.params;
}

// ---------------------------------------------------------------------------------------------------
// The last linear layer and the closed form A.2 in CONTRACTED form.  A.2 needs the Jacobian J = W_{L+1} T_L only
// through two contractions, so J (d x d nb per sample) is never formed:
//     R_n[m]  = sum_{j>=n} s_j T_L[j][m]                       (suffix sums over the tangent directions)
//     g_p     = W_{L+1}[p] . R_n      = sum_{j>=n} s_j J_{i,n,j}        p = (i, n), n <= min(i, nb-1)
//     D_p     = W_{L+1}[p] . T_L[i]   = J_{i,n,i}
//     Q_i     = sum_{n<=min(i,nb-1)} [ 2 f_p SX_n - 2 FX_n + g_p - SD_n ],      v_i = s_i Q_i
//     SX_n = sum_{j>=n} x_j,  FX_n = sum_{j>=n} f_{(j,n)} x_j,  SD_n = sum_{j>=n} s_j D_{(j,n)}
// which is the same sum as velocity_field.py:53-79 (the j = i terms cancel) at 3 dot products per kept output row
// instead of d + 1: the last layer costs O(3 R H) instead of O(H sum_n (d-n)^2) (13x fewer flops per evaluation at
// d=64, nb=63, H=256).  Reverse (A.3 + the last layer of A.4), with Qb_i = vbar_i s_i, SQ_n = sum_{i>=n} Qb_i:
//     fb_p = 2 (Qb_i SX_n - x_i SQ_n)     gb_p = Qb_i     Db_p = -s_i SQ_n
//     xbar_j = sum_{n<=min(j,nb-1)} [ SXb_n - 2 SQ_n f_{(j,n)} ],   SXb_n = sum_{i>=n} 2 f_{(i,n)} Qb_i
//     Rb_n = sum_{i>=n} Qb_i W[(i,n)]     CR_n = sum_{n'<=n} Rb_n'
//     hb_L = sum_p fb_p W[p]              Tb_L[j] = s_j CR_{min(j,nb-1)} + sum_n Db_{(j,n)} W[(j,n)]
//     sb_j = vbar_j Q_j - sum_n SQ_n D_{(j,n)} + CR_{min(j,nb-1)} . T_L[j];   xbar_j += 2 x_j sb_j
//     dW[p] += fb_p h_L + Qb_i R_n + Db_p T_L[i];   db[p] += fb_p
// ---------------------------------------------------------------------------------------------------
template <typename T>
struct Generic {
  static constexpr int JC = 8;  // tangent directions per register chunk
  static constexpr int KT = 4;  // output rows per weight-gradient tile
  static constexpr int NU = 4;  // last-layer rows per pass (independent loads in flight)

  template <typename S> GPF_DEV static T eps_part(const S& a);

  // ---- forward of one hidden linear layer li (0..L-1) + activation, lanes over output units ----
  // in: li == 0 -> the point x (S[d]) and time t; else hidden li (h, T)
  template <typename S>
  GPF_DEV static void forward_layer(const GArgs<T>& a, S* sc, int li, T t, int lane) {
    const GDims& g = a.g;
    const GLayout& l = a.lay;
    const int nin = g.n[li], no = g.n[li + 1], d = g.d;
    const T* __restrict__ Wt = a.params_t + g.woff[li];
    const T* __restrict__ bias = gparams(a) + g.boff[li];
    if (li == 0) {
      const S* x = sc + l.x;
      for (int k = lane; k < no; k += 32) {
        S acc = S(bias[k] + Wt[d * no + k] * t);
        for (int j = 0; j < d; ++j) acc = acc + Wt[j * no + k] * x[j];
        const S h = act_tanh(acc);
        sc[l.h[1] + k] = h;
        const S sp = T(1) - h * h;
        sc[l.sp[1] + k] = sp;
        for (int j = 0; j < d; ++j) {
          sc[l.U[1] + j * no + k] = S(Wt[j * no + k]);
          sc[l.T[1] + j * no + k] = sp * Wt[j * no + k];
        }
      }
      return;
    }
    const S* __restrict__ hin = sc + l.h[li];
    const S* __restrict__ Tin = sc + l.T[li];
    if constexpr (std::is_same<S, float>::value && GPF_GENERIC_MMA) {
      // tensor-core form (3xTF32): raw pre-activations / tangents first, activation pass after
      S* y0 = sc + l.h[li + 1];
      S* Y = sc + l.U[li + 1];
      if (a.on_chip) gm_rows_times_w<false>(hin, Tin, nin, d, nin, no, gparams(a) + g.woff[li], nin, bias, y0, Y, no, lane);
      else if (a.w_split != nullptr && nin % 8 == 0 && no % 64 == 0 && g.woff[li] % 4 == 0)
        gm_rows_times_w_fast(hin, Tin, nin, d, nin, no, a.w_split + 2 * (size_t)g.Ps + g.woff[li],
                             a.w_split + 3 * (size_t)g.Ps + g.woff[li], no, bias, y0, Y, no, lane);
      else gm_rows_times_w<true>(hin, Tin, nin, d, nin, no, Wt, no, bias, y0, Y, no, lane);
      for (int k = lane; k < no; k += 32) {
        const S h = act_tanh(y0[k]);
        y0[k] = h;
        const S sp = T(1) - h * h;
        sc[l.sp[li + 1] + k] = sp;
        for (int j = 0; j < d; ++j) sc[l.T[li + 1] + j * no + k] = sp * Y[j * no + k];
      }
      return;
    }
    for (int k = lane; k < no; k += 32) {
      S acc = S(bias[k]);
      for (int m = 0; m < nin; ++m) acc = acc + Wt[m * no + k] * hin[m];
      const S h = act_tanh(acc);
      sc[l.h[li + 1] + k] = h;
      const S sp = T(1) - h * h;
      sc[l.sp[li + 1] + k] = sp;
      for (int j0 = 0; j0 < d; j0 += JC) {
        S u[JC];
#pragma unroll
        for (int q = 0; q < JC; ++q) u[q] = S(T(0));
        for (int m = 0; m < nin; ++m) {
          const T wv = Wt[m * no + k];
#pragma unroll
          for (int q = 0; q < JC; ++q)
            if (j0 + q < d) u[q] = u[q] + wv * Tin[(j0 + q) * nin + m];
        }
#pragma unroll
        for (int q = 0; q < JC; ++q)
          if (j0 + q < d) {
            sc[l.U[li + 1] + (j0 + q) * no + k] = u[q];
            sc[l.T[li + 1] + (j0 + q) * no + k] = sp * u[q];
          }
      }
    }
  }

  // ---- last linear layer + A.2 (contracted form, see above): f, g, D, SX, Q, v from h_L, T_L, x ----
  template <typename S>
  GPF_DEV static void last_layer_forward(const GArgs<T>& a, S* sc, T bnd, int lane) {
    const GDims& g = a.g;
    const GLayout& l = a.lay;
    const int d = g.d, nb = g.nb, H = g.n[g.L];
    const T* __restrict__ W = gparams(a) + g.woff[g.L];  // [d nb][H]
    const T* __restrict__ bias = gparams(a) + g.boff[g.L];
    const S* x = sc + l.x;
    const S* __restrict__ hL = sc + l.h[g.L];
    const S* __restrict__ TL = sc + l.T[g.L];
    for (int j = lane; j < d; j += 32) sc[l.s + j] = x[j] * x[j] - bnd;
    __syncwarp();
    for (int m = lane; m < H; m += 32) {
      S acc = S(T(0));
      for (int j = d - 1; j >= 0; --j) {
        acc = acc + sc[l.s + j] * TL[j * H + m];
        if (j < nb) sc[l.R + j * H + m] = acc;
      }
    }
    __syncwarp();
    bool dots_done = false;
    if constexpr (std::is_same<S, float>::value) {
      if (a.vec4) {
        if (H <= 64) gl_dots_v<2>(W, bias, hL, TL, sc + l.R, sc + l.f, sc + l.G, sc + l.D, d, nb, H, lane);
        else gl_dots_v<4>(W, bias, hL, TL, sc + l.R, sc + l.f, sc + l.G, sc + l.D, d, nb, H, lane);
        dots_done = true;
      }
    }
    // three dot products per kept row, NU rows per pass so that 2 NU + 2 independent loads are in flight per lane
    for (int i = 0; i < d && !dots_done; ++i) {
      const int nmax = i < nb - 1 ? i : nb - 1;
      for (int n0 = 0; n0 <= nmax; n0 += NU) {
        S af[NU], ag[NU], ad[NU];
#pragma unroll
        for (int q = 0; q < NU; ++q) af[q] = ag[q] = ad[q] = S(T(0));
        for (int m = lane; m < H; m += 32) {
          T w[NU];
          S r[NU];
#pragma unroll
          for (int q = 0; q < NU; ++q) {
            const int nq = n0 + q <= nmax ? n0 + q : nmax;  // clamped rows are discarded below
            w[q] = W[(i * nb + nq) * H + m];
            r[q] = sc[l.R + nq * H + m];
          }
          const S hm = hL[m], ti = TL[i * H + m];
#pragma unroll
          for (int q = 0; q < NU; ++q) {
            af[q] = af[q] + w[q] * hm;
            ag[q] = ag[q] + w[q] * r[q];
            ad[q] = ad[q] + w[q] * ti;
          }
        }
        static_assert(NU == 4, "warp_sum4");
        const S rf = warp_sum4(af, lane), rg = warp_sum4(ag, lane), rd = warp_sum4(ad, lane);
        const int q = 2 * ((lane >> 4) & 1) + ((lane >> 3) & 1);  // the row whose totals this lane holds
        if ((lane & 7) == 0 && n0 + q <= nmax) {
          const int p = i * nb + n0 + q;
          sc[l.f + p] = rf + bias[p];
          sc[l.G + p] = rg;
          sc[l.D + p] = rd;
        }
      }
    }
    __syncwarp();
    for (int n = lane; n < nb; n += 32) {
      S sx = S(T(0)), fx = S(T(0)), sd = S(T(0));
      for (int j = n; j < d; ++j) {
        sx = sx + x[j];
        fx = fx + sc[l.f + j * nb + n] * x[j];
        sd = sd + sc[l.s + j] * sc[l.D + j * nb + n];
      }
      sc[l.SX + n] = sx;
      sc[l.CN + n] = T(-2) * fx - sd;
    }
    __syncwarp();
    for (int i = lane; i < d; i += 32) {
      S Q = S(T(0));
      const int nmax = i < nb - 1 ? i : nb - 1;
      for (int n = 0; n <= nmax; ++n)
        Q = Q + (T(2) * sc[l.f + i * nb + n]) * sc[l.SX + n] + sc[l.G + i * nb + n] + sc[l.CN + n];
      sc[l.Q + i] = Q;
      sc[l.v + i] = sc[l.s + i] * Q;
    }
    __syncwarp();
  }

  // velocity at the point stored in sc[l.x], result in sc[l.v]; keeps the activations for a following reverse
  template <typename S>
  __device__ __noinline__ static void eval_forward(const GArgs<T>& a, S* sc, T t, T bnd, int lane) {
    for (int li = 0; li < a.g.L; ++li) {
      forward_layer<S>(a, sc, li, t, lane);
      __syncwarp();
    }
    last_layer_forward<S>(a, sc, bnd, lane);
  }

  // ---- weight-gradient accumulation of hidden linear layer li < L: gacc[k][m] += eps/real part of
  //      ab[k] * in[m] + sum_j Ub[j][k] * Tin[j][m];  lanes over m, KT rows per tile ----
  template <typename S>
  GPF_DEV static void grad_layer(const GArgs<T>& a, const S* sc, T* gacc, int li, T t, const S* ab, const S* Ub, int lane) {
    const GDims& g = a.g;
    const GLayout& l = a.lay;
    const int nin = g.n[li], no = g.n[li + 1], d = g.d;
    T* __restrict__ gw = gacc + g.woff[li];
    T* __restrict__ gb = gacc + g.boff[li];
    // every element of the partial is always updated by the same lane: `red` keeps the sum order (see red_add_v2)
    auto acc_add = [&](T* p, T v) {
      if constexpr (std::is_same<T, float>::value && GPF_GENERIC_RED) {
        if (a.grad_red) { red_add_f32(p, v); return; }
      }
      *p += v;
    };
    for (int k = lane; k < no; k += 32) acc_add(gb + k, eps_part<S>(ab[k]));
    if (li == 0) {
      // W0[k][j] += ab[k] x_j + Ub[j][k];  W0[k][d] += ab[k] t      (lanes over k; rows are short)
      const S* x = sc + l.x;
      for (int k = lane; k < no; k += 32) {
        const S abk = ab[k];
        for (int j = 0; j < d; ++j) acc_add(gw + k * nin + j, eps_part<S>(abk * x[j] + Ub[j * no + k]));
        acc_add(gw + k * nin + d, eps_part<S>(abk * t));
      }
      return;
    }
    const S* __restrict__ hin = sc + l.h[li];
    const S* __restrict__ Tin = sc + l.T[li];
    if constexpr (std::is_same<S, float>::value && GPF_GENERIC_MMA) {
      gm_weight_grad(ab, Ub, no, hin, Tin, nin, d, no, nin, gw, lane, a.grad_red != 0);
      return;
    }
    for (int k0 = 0; k0 < no; k0 += KT) {
      for (int m = lane; m < nin; m += 32) {
        S acc[KT];
        const S hm = hin[m];
#pragma unroll
        for (int q = 0; q < KT; ++q) acc[q] = (k0 + q < no) ? ab[k0 + q] * hm : S(T(0));
        for (int j = 0; j < d; ++j) {
          const S tv = Tin[j * nin + m];
#pragma unroll
          for (int q = 0; q < KT; ++q)
            if (k0 + q < no) acc[q] = acc[q] + Ub[j * no + k0 + q] * tv;
        }
#pragma unroll
        for (int q = 0; q < KT; ++q)
          if (k0 + q < no) acc_add(gw + (k0 + q) * nin + m, eps_part<S>(acc[q]));
      }
    }
  }

  // ---- reverse through hidden linear layer 1 <= li < L: (ab[k], Ub[j][k]) -> hb[m], Tb[j][m] of hidden li ----
  template <typename S>
  GPF_DEV static void backprop_layer(const GArgs<T>& a, S* sc, int li, const S* ab, const S* Ub, int lane) {
    const GDims& g = a.g;
    const GLayout& l = a.lay;
    const int nin = g.n[li], no = g.n[li + 1], d = g.d;
    const T* __restrict__ W = gparams(a) + g.woff[li];
    if constexpr (std::is_same<S, float>::value && GPF_GENERIC_MMA) {
      if (!a.on_chip && a.w_split != nullptr && no % 8 == 0 && nin % 64 == 0 && g.woff[li] % 4 == 0)
        gm_rows_times_w_fast(ab, Ub, no, d, no, nin, a.w_split + g.woff[li], a.w_split + (size_t)g.Ps + g.woff[li], nin,
                             nullptr, sc + l.hb[li], sc + l.Tb[li], nin, lane);
      else gm_rows_times_w<true>(ab, Ub, no, d, no, nin, W, nin, nullptr, sc + l.hb[li], sc + l.Tb[li], nin, lane);
      return;
    }
    for (int m = lane; m < nin; m += 32) {
      S acc = S(T(0));
      for (int k = 0; k < no; ++k) acc = acc + W[k * nin + m] * ab[k];
      sc[l.hb[li] + m] = acc;
      for (int j0 = 0; j0 < d; j0 += JC) {
        S u[JC];
#pragma unroll
        for (int q = 0; q < JC; ++q) u[q] = S(T(0));
        for (int k = 0; k < no; ++k) {
          const T wv = W[k * nin + m];
#pragma unroll
          for (int q = 0; q < JC; ++q)
            if (j0 + q < d) u[q] = u[q] + wv * Ub[(j0 + q) * no + k];
        }
#pragma unroll
        for (int q = 0; q < JC; ++q)
          if (j0 + q < d) sc[l.Tb[li] + (j0 + q) * nin + m] = u[q];
      }
    }
  }

  // ---- reverse of last_layer_forward: cotangent sc[l.vbar] -> hb_L, Tb_L, first part of xbar, dW_{L+1}, db_{L+1} ----
  template <typename S>
  GPF_DEV static void last_layer_reverse(const GArgs<T>& a, S* sc, T* gacc, int lane) {
    const GDims& g = a.g;
    const GLayout& l = a.lay;
    const int d = g.d, nb = g.nb, no = g.nout, L = g.L, H = g.n[L];
    const T* __restrict__ W = gparams(a) + g.woff[L];
    T* __restrict__ gw = gacc + g.woff[L];
    T* __restrict__ gb = gacc + g.boff[L];
    const S* x = sc + l.x;
    const S* __restrict__ hL = sc + l.h[L];
    const S* __restrict__ TL = sc + l.T[L];
    for (int j = lane; j < d; j += 32) {
      sc[l.Qb + j] = sc[l.vbar + j] * sc[l.s + j];
      sc[l.sb + j] = sc[l.vbar + j] * sc[l.Q + j];
    }
    __syncwarp();
    for (int n = lane; n < nb; n += 32) {
      S sq = S(T(0)), sxb = S(T(0));
      for (int i = n; i < d; ++i) {
        const S qi = sc[l.Qb + i];
        sq = sq + qi;
        sxb = sxb + (T(2) * sc[l.f + i * nb + n]) * qi;
      }
      sc[l.SQ + n] = sq;
      sc[l.SXb + n] = sxb;
    }
    __syncwarp();
    for (int p = lane; p < no; p += 32) {
      const int i = p / nb, n = p % nb;
      S fbp = S(T(0));
      if (n <= i) {
        fbp = T(2) * (sc[l.Qb + i] * sc[l.SX + n] - x[i] * sc[l.SQ + n]);
        gb[p] += eps_part<S>(fbp);
      }
      sc[l.fb + p] = fbp;
    }
    for (int j = lane; j < d; j += 32) {
      S xb = S(T(0)), sbj = sc[l.sb + j];
      const int nmax = j < nb - 1 ? j : nb - 1;
      for (int n = 0; n <= nmax; ++n) {
        const S sq = sc[l.SQ + n];
        xb = xb + sc[l.SXb + n] - (T(2) * sq) * sc[l.f + j * nb + n];
        sbj = sbj - sq * sc[l.D + j * nb + n];
      }
      sc[l.xbar + j] = xb;
      sc[l.sb + j] = sbj;
    }
    __syncwarp();
    bool passes_done = false;
    if constexpr (std::is_same<S, float>::value) {
      if (a.vec4) {
        if (H <= 64)
          gl_reverse_v<2>(W, gw, hL, TL, sc + l.R, sc + l.fb, sc + l.Qb, sc + l.s, sc + l.SQ, sc + l.hb[L], sc + l.Tb[L],
                          sc + l.sb, d, nb, H, lane, a.grad_red != 0);
        else
          gl_reverse_v<4>(W, gw, hL, TL, sc + l.R, sc + l.fb, sc + l.Qb, sc + l.s, sc + l.SQ, sc + l.hb[L], sc + l.Tb[L],
                          sc + l.sb, d, nb, H, lane, a.grad_red != 0);
        passes_done = true;
      }
    }
    for (int m0 = 0; m0 < H && !passes_done; m0 += 32) {
      const int m = m0 + lane;
      const bool act = m < H;
      S hbm = S(T(0));
      const S hm = act ? hL[m] : S(T(0));
      // pass A: rows in storage order; hb_L, the D part of Tb_L, dW
      for (int i = 0; i < d; ++i) {
        const int nmax = i < nb - 1 ? i : nb - 1;
        const S Ti = act ? TL[i * H + m] : S(T(0));
        const S qbi = sc[l.Qb + i];
        const S nsi = -sc[l.s + i];
        S Ei = S(T(0));
        if (act)
          for (int n0 = 0; n0 <= nmax; n0 += NU) {
            T w[NU], go[NU];
            S r[NU], fbp[NU], dbp[NU];
#pragma unroll
            for (int q = 0; q < NU; ++q) {  // all loads of the pass first (the dW stores below may alias for the compiler)
              const int nq = n0 + q <= nmax ? n0 + q : nmax;
              const int p = i * nb + nq;
              w[q] = W[p * H + m];
              go[q] = gw[p * H + m];
              r[q] = sc[l.R + nq * H + m];
              fbp[q] = sc[l.fb + p];
              dbp[q] = nsi * sc[l.SQ + nq];
            }
#pragma unroll
            for (int q = 0; q < NU; ++q)
              if (n0 + q <= nmax) {
                hbm = hbm + fbp[q] * w[q];
                Ei = Ei + dbp[q] * w[q];
                gw[(i * nb + n0 + q) * H + m] = go[q] + eps_part<S>(fbp[q] * hm + qbi * r[q] + dbp[q] * Ti);
              }
          }
        if (act) sc[l.Tb[L] + i * H + m] = Ei;
      }
      if (act) sc[l.hb[L] + m] = hbm;
      // pass B: blocks n in order; Rb_n, its running sum CR_n, the R part of Tb_L and of sb
      S CR = S(T(0));
      for (int n = 0; n < nb; ++n) {
        S rb = S(T(0));
        if (act) {
          int i = n;
          for (; i + NU <= d; i += NU) {
            T w[NU];
#pragma unroll
            for (int q = 0; q < NU; ++q) w[q] = W[((i + q) * nb + n) * H + m];
#pragma unroll
            for (int q = 0; q < NU; ++q) rb = rb + sc[l.Qb + i + q] * w[q];
          }
          for (; i < d; ++i) rb = rb + sc[l.Qb + i] * W[(i * nb + n) * H + m];
        }
        CR = CR + rb;
        const int jhi = (n == nb - 1) ? d - 1 : n;
        for (int j = n; j <= jhi; ++j) {
          S part = S(T(0));
          if (act) {
            part = CR * TL[j * H + m];
            sc[l.Tb[L] + j * H + m] = sc[l.Tb[L] + j * H + m] + sc[l.s + j] * CR;
          }
          part = warp_sum(part);
          if (lane == 0) sc[l.sb + j] = sc[l.sb + j] + part;
        }
      }
    }
    __syncwarp();
    for (int j = lane; j < d; j += 32) sc[l.xbar + j] = sc[l.xbar + j] + (T(2) * x[j]) * sc[l.sb + j];
    __syncwarp();
  }

  // ---- full reverse sweep of one velocity evaluation (A.3 + A.4).  Needs the activations of eval_forward at
  //      sc[l.x]; cotangent in sc[l.vbar]; result J_v^T vbar in sc[l.xbar]; parameter gradients into gacc ----
  template <typename S>
  __device__ __noinline__ static void eval_reverse(const GArgs<T>& a, S* sc, T* gacc, T t, int lane) {
    const GDims& g = a.g;
    const GLayout& l = a.lay;
    const int d = g.d;
    last_layer_reverse<S>(a, sc, gacc, lane);
    const S* ab = sc + l.hb[g.L];
    const S* Ub = sc + l.Tb[g.L];
    for (int li = g.L; li >= 1; --li) {
      if (li < g.L) {
        grad_layer<S>(a, sc, gacc, li, t, ab, Ub, lane);
        backprop_layer<S>(a, sc, li, ab, Ub, lane);
        __syncwarp();
      }
      // through the activation of hidden li (in place: hb -> ab, Tb -> Ub)
      const int H = g.n[li];
      for (int k = lane; k < H; k += 32) {
        const S h = sc[l.h[li] + k], sp = sc[l.sp[li] + k];
        S rs = S(T(0));
        for (int j = 0; j < d; ++j) rs = rs + sc[l.Tb[li] + j * H + k] * sc[l.U[li] + j * H + k];
        const S spp = (T(-2) * h) * sp;
        sc[l.hb[li] + k] = sc[l.hb[li] + k] * sp + rs * spp;
        for (int j = 0; j < d; ++j) sc[l.Tb[li] + j * H + k] = sp * sc[l.Tb[li] + j * H + k];
      }
      __syncwarp();
      ab = sc + l.hb[li];
      Ub = sc + l.Tb[li];
    }
    grad_layer<S>(a, sc, gacc, 0, t, ab, Ub, lane);
    {  // xbar_j += sum_k W0[k][j] ab_1[k]
      const int H = g.n[1], nin = g.n[0];
      const T* __restrict__ W0 = gparams(a) + g.woff[0];
      for (int j = lane; j < d; j += 32) {
        S acc = sc[l.xbar + j];
        for (int k = 0; k < H; ++k) acc = acc + W0[k * nin + j] * ab[k];
        sc[l.xbar + j] = acc;
      }
    }
    __syncwarp();
  }
};

template <typename T> template <typename S>
GPF_DEV T Generic<T>::eps_part(const S& a) { return comp<IsDual<S>::value ? 1 : 0>(a); }

// per-warp real-valued work vectors (elements of T), after the scalar scratch
struct GVec { int xcur, ys, ksum, xb, ybsum, ybprev, kb, ystage, traj, aw, az, ub, stg, total; };
inline GVec make_gvec(const GDims& g, int n_steps) {
  GVec v;
  int off = 0;
  auto take = [&](int n) { int o = off; off += (n + 3) / 4 * 4; return o; };
  v.xcur = take(g.d); v.ys = take(g.d); v.ksum = take(g.d); v.xb = take(g.d); v.ybsum = take(g.d);
  v.ybprev = take(g.d); v.kb = take(g.d); v.ystage = take(4 * g.d); v.traj = take((n_steps + 1) * g.d);
  v.aw = take(g.d); v.az = take(g.d); v.ub = take(g.d);
  v.stg = take(n_steps * 4 * g.d);   // RK4 stage points of every step (kept for the reverse sweep)
  v.total = off;
  return v;
}

template <typename T>
struct GKernelArgs {
  GArgs<T> a;
  GVec vec;
  size_t vec_offset_bytes;  // offset of the T work vectors inside one warp's scratch
};

// epsilon part of a(x) = (v(x + delta v(x,t), t+delta) - v(x,t)) / delta along `dir` (A.7), result in out[d]
template <typename T>
__device__ __noinline__ void g_pen_tangent(const GArgs<T>& a, char* wscr, const T* xpt, const T* dir, T t, T bnd, T delta, T* out,
                           int lane) {
  typedef Dual<T> DS;
  DS* sc = reinterpret_cast<DS*>(wscr);
  const GLayout& l = a.lay;
  const int d = a.g.d;
  for (int j = lane; j < d; j += 32) sc[l.x + j] = DS(xpt[j], dir[j]);
  __syncwarp();
  Generic<T>::template eval_forward<DS>(a, sc, t, bnd, lane);
  for (int j = lane; j < d; j += 32) {
    const DS v1 = sc[l.v + j];
    out[j] = -v1.e;
    sc[l.x + j] = sc[l.x + j] + delta * v1;
  }
  __syncwarp();
  Generic<T>::template eval_forward<DS>(a, sc, t + delta, bnd, lane);
  for (int j = lane; j < d; j += 32) out[j] = (out[j] + sc[l.v + j].e) / delta;
  __syncwarp();
}

// reverse of g_pen_tangent for one direction; ubar = cotangent of the epsilon part; accumulates into xb[d]
template <typename T>
__device__ __noinline__ void g_pen_reverse(const GArgs<T>& a, char* wscr, T* gacc, const T* xpt, const T* dir, const T* ubar, T t, T bnd,
                           T delta, T* xb, T* tmp /* d, T */, int lane) {
  typedef Dual<T> DS;
  DS* sc = reinterpret_cast<DS*>(wscr);
  const GLayout& l = a.lay;
  const int d = a.g.d;
  // forward 1 to obtain X2 = X + delta v1
  for (int j = lane; j < d; j += 32) sc[l.x + j] = DS(xpt[j], dir[j]);
  __syncwarp();
  Generic<T>::template eval_forward<DS>(a, sc, t, bnd, lane);
  for (int j = lane; j < d; j += 32) sc[l.x + j] = sc[l.x + j] + delta * sc[l.v + j];
  __syncwarp();
  // eval 2 forward (keep) + reverse with V2b = (ubar/delta, 0)
  Generic<T>::template eval_forward<DS>(a, sc, t + delta, bnd, lane);
  for (int j = lane; j < d; j += 32) sc[l.vbar + j] = DS(ubar[j] / delta, T(0));
  __syncwarp();
  Generic<T>::template eval_reverse<DS>(a, sc, gacc, t + delta, lane);
  // X2b -> xb (eps part) and V1b = -ubar/delta + delta X2b ; stash V1b in vbar after recomputing eval 1
  DS* x2b = sc + l.Qb;  // Qb is free after eval_reverse: reuse as a d-vector stash
  for (int j = lane; j < d; j += 32) {
    const DS v = sc[l.xbar + j];
    x2b[j] = v;
    xb[j] += v.e;
  }
  __syncwarp();
  for (int j = lane; j < d; j += 32) {
    tmp[j] = x2b[j].v;          // stash both parts in T vectors across the forward recompute
    tmp[d + j] = x2b[j].e;
    sc[l.x + j] = DS(xpt[j], dir[j]);
  }
  __syncwarp();
  Generic<T>::template eval_forward<DS>(a, sc, t, bnd, lane);
  for (int j = lane; j < d; j += 32)
    sc[l.vbar + j] = DS(-ubar[j] / delta, T(0)) + delta * DS(tmp[j], tmp[d + j]);
  __syncwarp();
  Generic<T>::template eval_reverse<DS>(a, sc, gacc, t, lane);
  for (int j = lane; j < d; j += 32) xb[j] += sc[l.xbar + j].e;
  __syncwarp();
}

// this warp's scratch; in on-chip mode also stages the flat parameters into the CTA's shared memory (all threads of
// the CTA must call it)
template <typename T>
GPF_DEV char* generic_warp_scratch(const GArgs<T>& a, long long gwarp) {
  extern __shared__ __align__(16) unsigned char g_dyn_smem[];
  if (!a.on_chip) return a.scratch + (size_t)gwarp * a.scratch_stride;
  T* ps = reinterpret_cast<T*>(g_dyn_smem);
  for (int i = threadIdx.x; i < a.g.P; i += blockDim.x) ps[i] = a.params[i];
  __syncthreads();
  return reinterpret_cast<char*>(g_dyn_smem) + a.smem_scratch_off + (size_t)(threadIdx.x >> 5) * a.scratch_stride;
}

template <typename T, int MODE, bool EULER>
__global__ void __launch_bounds__(kGenericMaxThreads, 1) gpflow_generic_kernel(const GKernelArgs<T> ka) {
  const GArgs<T>& a = ka.a;
  const GLayout& l = a.lay;
  const GVec& vv = ka.vec;
  const int lane = threadIdx.x & 31;
  const long long gwarp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  const int d = a.g.d;
  char* wscr = generic_warp_scratch<T>(a, gwarp);
  T* sc = reinterpret_cast<T*>(wscr);
  T* vec = reinterpret_cast<T*>(wscr + ka.vec_offset_bytes);
  T* part = a.partials + (size_t)gwarp * a.npart;
  T* gacc = part + 3;
  if (MODE != MODE_FWD)
    for (int i = lane; i < a.g.P; i += 32) gacc[i] = T(0);
  __syncwarp();

  const bool wrap = (a.flags & 0x1u) != 0;
  const bool want_lagr = (a.flags & 0x2u) != 0;
  const T bnd = T(1) - Eps<T>::value;
  const T dt = a.dt, hdt = a.dt / T(2), c6 = a.dt / T(6);
  const T delta = T(2.0 * sqrt((double)Eps<T>::value));
  const T sq2s = T(1.4142135623730951) * a.sigma;
  T reg_cot = T(1);
  if (MODE == MODE_BWD && a.reg_cotangent != nullptr) reg_cot = *a.reg_cotangent;
  const T pen_w = a.euler_coef * a.loss_scale;
  T ot_acc = T(0), pen_acc = T(0), lagr_acc = T(0);  // lane-partial sums

  T* xcur = vec + vv.xcur; T* ys = vec + vv.ys; T* ksum = vec + vv.ksum; T* xb = vec + vv.xb;
  T* ybsum = vec + vv.ybsum; T* ybprev = vec + vv.ybprev; T* pen_tmp = vec + vv.ystage; T* traj = vec + vv.traj;
  T* aw = vec + vv.aw; T* az = vec + vv.az; T* ub = vec + vv.ub; T* stg = vec + vv.stg;

  const int wic = threadIdx.x >> 5, wpc = blockDim.x >> 5;
  for (long long bb = (long long)blockIdx.x * wpc; bb < a.B; bb += nwarps) {
    if (bb + wic >= a.B) break;
    const long long b = bb + wic;
    // The CTA's warps run the same sequence of evaluations on different samples.  A named barrier over the warps that have
    // a sample in this round keeps them within one evaluation of each other, so that they share the instruction-cache
    // footprint and the weight lines in L1 (ncu at C3: stall_no_inst 27 % with free-running warps).
    const long long left = a.B - bb;
    const unsigned n_act_threads = 32u * (unsigned)(left < wpc ? left : wpc);
    auto cta_align = [&]() {
      if (a.cta_align && n_act_threads > 32u) asm volatile("bar.sync 1, %0;" ::"r"(n_act_threads) : "memory");
    };
    for (int j = lane; j < d; j += 32) {
      const T xi = a.x_in[b * d + j];
      xcur[j] = wrap ? erf_real(xi / sq2s) : xi;
    }
    __syncwarp();
    // ---------------- forward time loop ----------------
    T t = a.t0;
    for (int i = 0; i < a.n_steps; ++i) {
      for (int j = lane; j < d; j += 32) {
        traj[i * d + j] = xcur[j];
        if (a.traj_out != nullptr) a.traj_out[((long long)i * a.B + b) * d + j] = xcur[j];
        ys[j] = xcur[j];
      }
      __syncwarp();
      for (int sg = 0; sg < 4; ++sg) {
        const T ts = sg == 0 ? t : (sg == 3 ? t + dt : t + hdt);
        for (int j = lane; j < d; j += 32) {
          sc[l.x + j] = ys[j];
          if (MODE != MODE_FWD) stg[(i * 4 + sg) * d + j] = ys[j];
        }
        __syncwarp();
        cta_align();
        Generic<T>::template eval_forward<T>(a, sc, ts, bnd, lane);
        const T wk = (sg == 1 || sg == 2) ? T(2) : T(1);
        const T ak = sg == 2 ? dt : hdt;
        for (int j = lane; j < d; j += 32) {
          const T kv = sc[l.v + j];
          ksum[j] = sg == 0 ? kv : ksum[j] + wk * kv;
          ys[j] = xcur[j] + ak * kv;
        }
        __syncwarp();
      }
      for (int j = lane; j < d; j += 32) xcur[j] = xcur[j] + c6 * ksum[j];
      __syncwarp();
      if (want_lagr) {
        for (int j = lane; j < d; j += 32) sc[l.x + j] = xcur[j];
        __syncwarp();
        Generic<T>::template eval_forward<T>(a, sc, t, bnd, lane);
        for (int j = lane; j < d; j += 32) lagr_acc += T(0.5) * sc[l.v + j] * sc[l.v + j];
        __syncwarp();
      }
      if (EULER && MODE == MODE_FWD) {
        const T* zv = a.zw + (size_t)(i * 2 + 0) * d;
        const T* wv = a.zw + (size_t)(i * 2 + 1) * d;
        cta_align();
        g_pen_tangent<T>(a, wscr, xcur, wv, t, bnd, delta, aw, lane);
        cta_align();
        g_pen_tangent<T>(a, wscr, xcur, zv, t, bnd, delta, az, lane);
        T q = T(0);
        for (int j = lane; j < d; j += 32) q += zv[j] * aw[j] - wv[j] * az[j];
        q = warp_sum(q);
        if (lane == 0) pen_acc += q * q;
      }
      t = t + dt;
    }
    for (int j = lane; j < d; j += 32) {
      traj[a.n_steps * d + j] = xcur[j];
      if (a.traj_out != nullptr) a.traj_out[((long long)a.n_steps * a.B + b) * d + j] = xcur[j];
    }
    // ---------------- wrapper epilogue ----------------
    T ny = T(0), nx = T(0);
    for (int j = lane; j < d; j += 32) {
      const T gT = xcur[j];
      T yv = gT;
      bool inband = true;
      if (wrap) {
        const T lim = T(1) - Eps<T>::value;
        const T gc = gT < -lim ? -lim : (gT > lim ? lim : gT);
        inband = (gT >= -lim) && (gT <= lim);
        yv = sq2s * erfinv_real(gc);
      }
      const T xi = a.x_in[b * d + j];
      ny += yv * yv;
      nx += xi * xi;
      if (a.y_out != nullptr) a.y_out[b * d + j] = yv;
      if (MODE != MODE_FWD) {
        T yb;
        if (MODE == MODE_TRAIN) {
          const T diff = a.x_data[b * d + j] - yv;
          ot_acc += diff * diff;
          yb = T(-2) * diff * a.loss_scale;
        } else {
          yb = a.grad_y[b * d + j];
        }
        if (wrap) {
          const T dydg = T(1.2533141373155001) * a.sigma * exp_real(yv * yv / (T(2) * a.sigma * a.sigma));
          xb[j] = inband ? yb * dydg : T(0);
        } else {
          xb[j] = yb;
        }
      }
    }
    if (a.logdet_out != nullptr) {
      ny = warp_sum(ny);
      nx = warp_sum(nx);
      if (lane == 0) a.logdet_out[b] = (ny - nx) / (T(2) * a.sigma * a.sigma);
    }
    __syncwarp();
    if (MODE == MODE_FWD) continue;

    // ---------------- reverse time loop ----------------
    // times: t_i by repeated addition -> recompute the forward sequence (cheap) for each i
    for (int i = a.n_steps - 1; i >= 0; --i) {
      T ti = a.t0;
      for (int q = 0; q < i; ++q) ti = ti + dt;
      if (EULER) {
        const T* xn = traj + (size_t)(i + 1) * d;
        const T* zv = a.zw + (size_t)(i * 2 + 0) * d;
        const T* wv = a.zw + (size_t)(i * 2 + 1) * d;
        cta_align();
        g_pen_tangent<T>(a, wscr, xn, wv, ti, bnd, delta, aw, lane);
        cta_align();
        g_pen_tangent<T>(a, wscr, xn, zv, ti, bnd, delta, az, lane);
        T q = T(0);
        for (int j = lane; j < d; j += 32) q += zv[j] * aw[j] - wv[j] * az[j];
        q = warp_sum(q);
        if (lane == 0) pen_acc += q * q;
        const T gq = T(2) * q * pen_w * reg_cot;
        for (int j = lane; j < d; j += 32) ub[j] = gq * zv[j];
        __syncwarp();
        cta_align();
        g_pen_reverse<T>(a, wscr, gacc, xn, wv, ub, ti, bnd, delta, xb, pen_tmp /* scratch 2d */, lane);
        for (int j = lane; j < d; j += 32) ub[j] = -gq * wv[j];
        __syncwarp();
        cta_align();
        g_pen_reverse<T>(a, wscr, gacc, xn, zv, ub, ti, bnd, delta, xb, pen_tmp, lane);
      }
      // the four RK4 stage points of step i were kept by the forward sweep (no re-integration of the step)
      const T* ystage = stg + (size_t)i * 4 * d;
      for (int j = lane; j < d; j += 32) { ybsum[j] = T(0); ybprev[j] = T(0); }
      __syncwarp();
      for (int sg = 3; sg >= 0; --sg) {
        const T ts = sg == 0 ? ti : (sg == 3 ? ti + dt : ti + hdt);
        const T wk = (sg == 1 || sg == 2) ? T(2) * c6 : c6;
        const T ak = sg == 3 ? T(0) : (sg == 2 ? dt : hdt);
        for (int j = lane; j < d; j += 32) sc[l.x + j] = ystage[sg * d + j];
        __syncwarp();
        cta_align();
        Generic<T>::template eval_forward<T>(a, sc, ts, bnd, lane);
        for (int j = lane; j < d; j += 32) sc[l.vbar + j] = wk * xb[j] + ak * ybprev[j];
        __syncwarp();
        Generic<T>::template eval_reverse<T>(a, sc, gacc, ts, lane);
        for (int j = lane; j < d; j += 32) {
          const T yb = sc[l.xbar + j];
          ybprev[j] = yb;
          ybsum[j] += yb;
        }
        __syncwarp();
      }
      for (int j = lane; j < d; j += 32) xb[j] = xb[j] + ybsum[j];
      __syncwarp();
    }
    if (a.grad_x_out != nullptr) {
      for (int j = lane; j < d; j += 32) {
        T gx = xb[j];
        if (wrap) {
          const T xi = a.x_in[b * d + j];
          gx *= T(0.7978845608028654) / a.sigma * exp_real(-xi * xi / (T(2) * a.sigma * a.sigma));
        }
        a.grad_x_out[b * d + j] = gx;
      }
    }
    __syncwarp();
  }
  ot_acc = warp_sum(ot_acc);
  pen_acc = warp_sum(pen_acc);
  lagr_acc = warp_sum(lagr_acc);
  if (lane == 0) {
    part[0] = ot_acc * a.loss_scale;
    part[1] = pen_acc * pen_w;
    part[2] = lagr_acc * a.dt_constraint * a.loss_scale;
  }
}

// VelocityField.__call__(x, t)  (velocity_field.py:118-126)
template <typename T>
__global__ void __launch_bounds__(kGenericMaxThreads, 1) gpflow_velocity_kernel(const GKernelArgs<T> ka) {
  const GArgs<T>& a = ka.a;
  const GLayout& l = a.lay;
  const int lane = threadIdx.x & 31;
  const long long gwarp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  T* sc = reinterpret_cast<T*>(generic_warp_scratch<T>(a, gwarp));
  const T bnd = T(1) - Eps<T>::value;
  for (long long b = gwarp; b < a.B; b += nwarps) {
    for (int j = lane; j < a.g.d; j += 32) sc[l.x + j] = a.x_in[b * a.g.d + j];
    __syncwarp();
    Generic<T>::template eval_forward<T>(a, sc, a.t_eval, bnd, lane);
    for (int j = lane; j < a.g.d; j += 32) a.v_out[b * a.g.d + j] = sc[l.v + j];
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------------------
// host side: workspace carving and launch
// ---------------------------------------------------------------------------------------------------
// per-warp partial vector [1 pad | ot, reg, lagr | P gradients | pad]: the gradients start 16-byte aligned
inline size_t generic_part_stride(int P) { return (size_t)(3 + P + 1 + 3) / 4 * 4; }

struct GenericPlan {
  long long nwarps;
  int warps_per_cta, on_chip;
  size_t smem_bytes, smem_scratch_off;
  size_t scratch_stride, vec_offset, partials_bytes, params_t_bytes, scratch_bytes, total;
};
// Two placements of the per-warp scratch (activations, tangents, adjoints of ONE velocity evaluation):
//   on-chip  when the flat parameters + at least 12 warps' scratch fit the CTA's shared memory: one CTA per SM of up
//            to 16 warps, every operand of the contractions at shared-memory latency (small latent shapes);
//   global   otherwise: 4 CTAs of 4 warps per SM, scratch in the workspace (L2 / HBM).  Measured at d=10, nb=9,
//            H=[50,50,50] (24 KB per warp, 7 warps fit): 45k samples/s on-chip vs 55k global -- 7 warps do not hide
//            the remaining latencies, so the threshold is 12 warps.  d=64, nb=63, H=256 needs 400 KB per warp.
inline GenericPlan generic_plan(const GpFlowDesc* d, int64_t B, int op, int sm_count) {
  const size_t es = d->dtype == GPFLOW_F64 ? 8 : 4;
  const GDims g = make_gdims(d);
  const GLayout l = make_glayout(g);
  const GVec v = make_gvec(g, d->n_steps);
  const bool dual = (d->flags & GPFLOW_FLAG_EULER_PEN) != 0;
  GenericPlan p;
  memset(&p, 0, sizeof(p));
  p.vec_offset = ((size_t)l.total * es * (dual ? 2 : 1) + 255) / 256 * 256;
  p.scratch_stride = (p.vec_offset + (size_t)v.total * es + 255) / 256 * 256;
  p.smem_scratch_off = ((size_t)g.P * es + 255) / 256 * 256;
  // global mode: 16 warps per SM as CTAs of GPFLOW_GENERIC_WPC warps
  static const int wpc_env = [] { const char* e = getenv("GPFLOW_GENERIC_WPC"); const int v = e ? atoi(e) : 0;
                                  return (v == 1 || v == 2 || v == 4 || v == 8 || v == 16) ? v : 0; }();
  // 16-warp CTAs where the warps are phase-aligned (small parameter sets, see launch_generic_impl); free-running warps do
  // slightly better as 4 CTAs of 4 warps per SM (C5: 3616 -> 3738 samples/s, same box)
  p.warps_per_cta = wpc_env ? wpc_env : ((size_t)g.P * es <= 256 * 1024 ? 16 : 4);
  long long nw = (long long)sm_count * 16;
  if (!wpc_env && B < nw) {   // small batch (one warp per sample): smaller CTAs so that the warps spread over all the SMs
    const long long per_sm = (B + sm_count - 1) / sm_count;
    while (p.warps_per_cta > 1 && p.warps_per_cta / 2 >= per_sm) p.warps_per_cta /= 2;
  }
  static const int on_chip_allowed = [] { const char* e = getenv("GPFLOW_GENERIC_ONCHIP"); return e == nullptr ? 1 : atoi(e); }();
  const int on_chip_min_warps = on_chip_allowed >= 2 ? on_chip_allowed : 12;   // GPFLOW_GENERIC_ONCHIP=<n >= 2>: experiment knob
  if (on_chip_allowed && (size_t)kGenericSmemBudget > p.smem_scratch_off + (size_t)on_chip_min_warps * p.scratch_stride) {
    int w = (int)(((size_t)kGenericSmemBudget - p.smem_scratch_off) / p.scratch_stride);
    if (w > kGenericMaxThreads / 32) w = kGenericMaxThreads / 32;
    p.on_chip = 1;
    p.warps_per_cta = w;
    p.smem_bytes = p.smem_scratch_off + (size_t)w * p.scratch_stride;
    nw = (long long)sm_count * w;
  }
  if (nw > B) nw = B;
  // bound the per-warp scratch + partial vectors (gradient accumulators) to 16 GiB of the 180 GB
  const size_t per_warp = (p.on_chip ? 0 : p.scratch_stride) + (size_t)(3 + (op == GPFLOW_OP_FORWARD ? 0 : g.P)) * es;
  const long long cap = (long long)((size_t)16 << 30) / (long long)per_warp;
  if (nw > cap) nw = cap;
  if (nw < 1) nw = 1;
  const long long nw16 = (nw + 15) / 16 * 16;
  nw = (nw + p.warps_per_cta - 1) / p.warps_per_cta * p.warps_per_cta;  // whole CTAs
  p.nwarps = nw;
  // sizes: global mode rounds to 16 warps whatever the CTA size, so that the device-independent bound (160 SMs) covers
  // every CTA size the real SM count may select
  if (!p.on_chip) nw = nw16;
  p.partials_bytes = (((size_t)nw * generic_part_stride(g.P) + 4) * es + 255) / 256 * 256;
  p.params_t_bytes = ((size_t)g.Ps * es * (es == 4 ? 5 : 1) + 255) / 256 * 256;   // Wt (+ the 4 tf32 split copies in fp32)
  p.scratch_bytes = p.on_chip ? 0 : (size_t)nw * p.scratch_stride;
  p.total = p.partials_bytes + p.params_t_bytes + p.scratch_bytes;
  return p;
}
inline size_t generic_pad_bytes(const GpFlowDesc* d, const GpFlowDesc* dp) {   // [padded params | index map]
  return ((size_t)make_gdims(dp).P * 4 + 255) / 256 * 256 + ((size_t)make_gdims(d).P * 4 + 255) / 256 * 256;
}
inline size_t generic_workspace(const GpFlowDesc* d, int64_t B, int op) {
  GpFlowDesc dp;
  if (generic_padded_desc(d, &dp)) return generic_plan(&dp, B, op, 160).total + generic_pad_bytes(d, &dp);
  return generic_plan(d, B, op, 160).total;  // device-independent upper bound (B200 has 148 SMs)
}

template <typename T, bool EULER>
int launch_generic_impl(const GpFlowDesc* d_user, int op, const T* params, const T* x_in, const T* x_data, const T* grad_y,
                   const T* reg_cot, const T* zw, T* y_out, T* logdet_out, T* traj_out, T* grad_x_out, T* grads_out,
                   T* scalars_out, int64_t B, void* ws, size_t wsb, cudaStream_t st, int sm_count, T* v_out,
                   double t_eval) {
  if (ws == nullptr || wsb < generic_workspace(d_user, B, op)) return GPFLOW_ERR_WORKSPACE;
  if (reinterpret_cast<uintptr_t>(ws) & 15) return GPFLOW_ERR_MISALIGNED;
  GpFlowDesc d_pad;
  const bool padded = !EULER && generic_padded_desc(d_user, &d_pad);
  const GpFlowDesc* d = padded ? &d_pad : d_user;
  const int* pad_map = nullptr;
  if (padded) {   // the padded parameters and the index map live at the END of the workspace
    const GDims g0 = make_gdims(d_user), gp = make_gdims(d);
    char* tail = reinterpret_cast<char*>(ws) + generic_plan(d, B, op, 160).total;
    T* pp = reinterpret_cast<T*>(tail);
    int* map = reinterpret_cast<int*>(tail + ((size_t)gp.P * 4 + 255) / 256 * 256);
    if (cudaMemsetAsync(pp, 0, (size_t)gp.P * sizeof(T), st) != cudaSuccess) { cudaGetLastError(); return GPFLOW_ERR_CUDA; }
    gpflow_pad_params_kernel<T><<<(g0.P + 255) / 256, 256, 0, st>>>(g0, gp, params, pp, map);
    if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
    params = pp;
    pad_map = map;
  }
  const GenericPlan pl = generic_plan(d, B, op, sm_count);
  GKernelArgs<T> ka;
  memset(&ka, 0, sizeof(ka));
  GArgs<T>& a = ka.a;
  a.g = make_gdims(d);
  a.lay = make_glayout(a.g);
  ka.vec = make_gvec(a.g, d->n_steps);
  ka.vec_offset_bytes = pl.vec_offset;
  char* w = reinterpret_cast<char*>(ws);
  a.partials = reinterpret_cast<T*>(w) + 1; w += pl.partials_bytes;   // + 1: see generic_part_stride
  T* pt = reinterpret_cast<T*>(w); w += pl.params_t_bytes;
  a.scratch = w;
  a.scratch_stride = pl.scratch_stride;
  a.params = params; a.params_t = pt; a.x_in = x_in; a.x_data = x_data; a.grad_y = grad_y; a.reg_cotangent = reg_cot;
  a.zw = zw; a.y_out = y_out; a.logdet_out = logdet_out; a.traj_out = traj_out; a.grad_x_out = grad_x_out;
  a.v_out = v_out; a.t_eval = (T)t_eval;
  a.B = B; a.n_steps = d->n_steps; a.npart = (int)generic_part_stride(a.g.P); a.flags = d->flags;
  a.vec4 = (sizeof(T) == 4 && a.g.n[a.g.L] % 4 == 0 && a.g.woff[a.g.L] % 4 == 0 &&
            (reinterpret_cast<uintptr_t>(params) & 15) == 0) ? 1 : 0;
  const bool rev = (d->flags & GPFLOW_FLAG_REVERSE) != 0;
  a.dt = rev ? -(T)d->dt : (T)d->dt;
  a.t0 = rev ? (T)(d->dt * (double)d->n_steps_nominal) : (T)0;
  a.dt_constraint = (T)d->dt_constraint; a.sigma = (T)d->sigma; a.euler_coef = (T)d->euler_coef;
  a.loss_scale = (T)d->loss_scale;
  a.w_split = sizeof(T) == 4 ? pt + a.g.Ps : nullptr;
  gpflow_transpose_params_kernel<T><<<(a.g.P + 255) / 256, 256, 0, st>>>(a.g, params, pt, const_cast<T*>(a.w_split));
  if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  a.on_chip = pl.on_chip;
  // phase alignment of the CTA's warps: measured +32 % at C3 (1.03e5 -> 1.36e5 samples/s, instruction-cache bound, 60 KB of
  // parameters) and -10 % at C5 (3260 -> 2920, 4.4 MB of parameters streamed per evaluation: free-running warps spread
  // that traffic) => on for parameter sets up to 256 KB.  GPFLOW_GENERIC_ALIGN=0/1 forces it (A/B measurements).
  static const int align_env = [] { const char* e = getenv("GPFLOW_GENERIC_ALIGN"); return e == nullptr ? -1 : (e[0] != '0'); }();
  static const int red_env = [] { const char* e = getenv("GPFLOW_GENERIC_RED"); return e == nullptr ? -1 : (e[0] != '0'); }();
  // same-box A/B: C3 1.39e5 -> 1.50e5 samples/s, C5 3357 -> 3496; GPFLOW_GENERIC_RED=0 restores the read-modify-write
  a.grad_red = red_env >= 0 ? red_env : 1;
  // The Euler instantiation (scalar dual arithmetic, an even larger kernel) gains more: C3 + Euler, B = 4096, same box,
  // alternating: 453 ms free-running, 144 - 147 ms aligned (9.0e3 -> 2.8e4 samples/s), results bit-identical.
  a.cta_align = align_env >= 0 ? align_env : ((size_t)a.g.P * sizeof(T) <= 256 * 1024 ? 1 : 0);
  a.smem_scratch_off = (unsigned)pl.smem_scratch_off;
  const unsigned grid = (unsigned)(pl.nwarps / pl.warps_per_cta), threads = 32u * (unsigned)pl.warps_per_cta;
  auto launch = [&](void (*kern)(const GKernelArgs<T>)) -> int {
    if (pl.smem_bytes > 48 * 1024 &&
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes) != cudaSuccess) {
      cudaGetLastError();
      return GPFLOW_ERR_CUDA;
    }
    kern<<<grid, threads, pl.smem_bytes, st>>>(ka);
    return cudaGetLastError() == cudaSuccess ? GPFLOW_OK : GPFLOW_ERR_CUDA;
  };
  if (v_out != nullptr) {
    if (EULER) return GPFLOW_ERR_BAD_DESC;
    return launch(gpflow_velocity_kernel<T>);
  }
  int rc;
  if (op == GPFLOW_OP_FORWARD) rc = launch(gpflow_generic_kernel<T, MODE_FWD, EULER>);
  else if (op == GPFLOW_OP_TRAIN_STEP) rc = launch(gpflow_generic_kernel<T, MODE_TRAIN, EULER>);
  else rc = launch(gpflow_generic_kernel<T, MODE_BWD, EULER>);
  if (rc != GPFLOW_OK) return rc;
  const int count = op == GPFLOW_OP_FORWARD ? 3 : 3 + make_gdims(d_user).P;
  if (scalars_out != nullptr || grads_out != nullptr) {
    if (padded)
      gpflow_finalize_mapped_kernel<T><<<(count + 127) / 128, 128, 0, st>>>(a.partials, (int)pl.nwarps, a.npart, count,
                                                                          pad_map, scalars_out, grads_out);
    else
      gpflow_finalize_kernel<T><<<(count + 127) / 128, 128, 0, st>>>(a.partials, (int)pl.nwarps, a.npart, count,
                                                                     scalars_out, grads_out);
    if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  }
  return GPFLOW_OK;
}

}  // namespace gpflow
