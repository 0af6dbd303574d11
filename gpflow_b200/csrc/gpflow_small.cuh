// gpflow_b200 -- "small-d" fused GP-flow kernels (sm_100a): one thread per sample, two hidden layers.
//
// Covers (reference file:line)                       SURVEY.md 8a row
//   VectorField.eval / .jacobian  vector_field.py:75-128         1   potentials + forward-mode tangents
//   divergence_free_vector_func   velocity_field.py:20-79        2   closed form A.2, pruned to (i>=n, j>=n)
//   RK4 compute_step / ode_solve  ode_order_method.py:66-71, ode_solver.py:70-81      3
//   ResNet.solve / constraints    res_net.py:199-317,360-364     4   Lagrangian reduction
//   Network.gp_flow               network.py:156-189             5   erf / clip / erfinv / log-det
//   Network.loss_gp_flow + backward network.py:70-84,104-111     6   OT cost, full reverse sweep, weight gradients
//   compute_euler_pen             res_net.py:319-358, velocity_field.py:180-185   7   (dual arithmetic)
//
// Layout: per-sample state (x, activations, tangents, adjoints) lives in registers for the whole time
// loop; the (tiny) MLP weights are staged once per CTA by a TMA bulk copy into shared memory and
// re-packed into 16-byte rows so that every weight read is a warp-uniform vector LDS.  The only cross-sample
// coupling is the weight gradient W_l += sum_b L_b (x) R_b, a [M x 32*(d+1)] x [32*(d+1) x N] contraction
// per warp and velocity evaluation: lanes stage their L/R rows in shared memory and each lane owns a 4x4
// register tile of the product (split over the sample axis when the layer has < 32 tiles); lane-private
// accumulators live in shared memory, are summed over lanes/warps once at the end of the kernel, written as
// one partial vector per CTA, and reduced across CTAs in a fixed order by gpflow_finalize (deterministic).
#pragma once
#include "gpflow_common.cuh"

namespace gpflow {

enum { MODE_FWD = 0, MODE_TRAIN = 1, MODE_BWD = 2 };

template <typename T>
struct KArgs {
  const T* params;         // [P]
  const T* x_in;           // [B,d]  x_map (wrap) or cube-space input
  const T* x_data;         // [B,d]  OT target (MODE_TRAIN)
  const T* grad_y;         // [B,d]  external cotangent (MODE_BWD)
  const T* reg_cotangent;  // device scalar or null
  const T* zw;             // [N,2,d]
  T* y_out;                // [B,d] nullable
  T* logdet_out;           // [B] nullable
  T* traj_out;             // [N+1,B,d] nullable
  T* grad_x_out;           // [B,d] nullable
  T* partials;             // [grid, npart]  : [3 scalars | P grads]
  T* traj_ws;              // [N+1, d, TS] trajectory scratch (TS = total threads)
  long long B;
  int n_steps;
  int npart;               // stride of `partials`
  unsigned flags;
  T dt;                    // signed solve dt (negative when reversed)
  T t0;
  T dt_constraint;
  T sigma;
  T euler_coef;
  T loss_scale;
};

// ---- 16-byte vector access on shared memory -------------------------------------------------
template <typename T> struct Vec16;
template <> struct Vec16<float> { typedef float4 type; static constexpr int N = 4; };
template <> struct Vec16<double> { typedef double2 type; static constexpr int N = 2; };

// Shared-memory row loads are issued as `asm volatile` vector loads: the packed weights never change after the
// prologue, and a plain C++ load lets the compiler prove that, hoist all ~330 weights out of the time loop and
// spill them.  Volatile keeps each load next to its use.
GPF_DEV void lds_vec(const float* p, float (&e)[4]) {
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(e[0]), "=f"(e[1]), "=f"(e[2]), "=f"(e[3]) : "r"(smem_u32(p)));
}
GPF_DEV void lds_vec(const double* p, double (&e)[2]) {
  asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(e[0]), "=d"(e[1]) : "r"(smem_u32(p)));
}
template <typename T, int N>
GPF_DEV void lds_row(const T* __restrict__ p, T (&out)[N]) {  // p 16-byte aligned, N multiple of Vec16<T>::N
  constexpr int VN = Vec16<T>::N;
#pragma unroll
  for (int i = 0; i < N / VN; ++i) {
    T e[VN];
    lds_vec(p + i * VN, e);
#pragma unroll
    for (int q = 0; q < VN; ++q) out[i * VN + q] = e[q];
  }
}
template <typename T, int N>
GPF_DEV void sts_row(T* __restrict__ p, const T (&in)[N]) {
  typedef typename Vec16<T>::type V;
  constexpr int VN = Vec16<T>::N;
#pragma unroll
  for (int i = 0; i < N / VN; ++i) {
    V v;
    T* e = reinterpret_cast<T*>(&v);
#pragma unroll
    for (int q = 0; q < VN; ++q) e[q] = in[i * VN + q];
    reinterpret_cast<V*>(p)[i] = v;
  }
}

__host__ __device__ constexpr int pow2_floor(int x) { int p = 1; while (p * 2 <= x) p *= 2; return p; }
// row stride (elements) >= n such that stride*sizeof(T)/16 is odd -> conflict-free 16-byte row accesses by 8 lanes
template <typename T> __host__ __device__ constexpr int odd16_stride(int n) {
  int vn = 16 / (int)sizeof(T);
  int s = round_up(n, vn);
  return ((s / vn) % 2 == 1) ? s : s + vn;
}

// ---- tiling of one weight-gradient contraction  C[M x NC] += sum_{s<32} L[s][M] (x) R[s][NC] -----
#ifndef GPF_V1_NOINLINE
#define GPF_V1_NOINLINE 1
#endif
// GPF_V1_ROLL: the two H2 x H1 contractions over the rows of W2 keep their row loop rolled (the row index then
// addresses h2 / U2 / ab2 / Ub2 dynamically, i.e. in local memory) -- trades a few LDL/STL for 15x less code
#ifndef GPF_V1_ROLL
#define GPF_V1_ROLL 0   // measured: 9.6e6 vs 1.49e7 samples/s on the Euler row (LDL/STL on the critical path), kept as an option
#endif
#if GPF_V1_ROLL
#define GPF_V1_ROLL_PRAGMA _Pragma("unroll 1")
#else
#define GPF_V1_ROLL_PRAGMA _Pragma("unroll")
#endif
#if GPF_V1_NOINLINE
#define GPF_V1_CALL __device__ __noinline__
#else
#define GPF_V1_CALL GPF_DEV
#endif

template <int M_, int NC_>
struct TileCfg {
  static constexpr int M = M_, NC = NC_;
  static constexpr int MP = round_up(M, 4), NCP = round_up(NC, 4);
  static constexpr int TM = MP / 4, TN = NCP / 4, NT = TM * TN;
  static constexpr int KS = NT >= 32 ? 1 : pow2_floor(32 / NT);    // split of the 32-sample axis
  static constexpr int TPL = NT >= 32 ? cdiv(NT, 32) : 1;          // tiles per lane
  static constexpr int ACC = TPL * 16 * 32;                         // lane-private accumulators per warp
};

template <typename T, int D_, int NB_, int H1_, int H2_, int BLOCK_>
struct SmallKernel {
  static constexpr int D = D_, NB = NB_, H1 = H1_, H2 = H2_, BLOCK = BLOCK_;
  static constexpr int NWARP = BLOCK / 32;
  static constexpr int NIN = D + 1, NOUT = D * NB;
  static constexpr int VN = Vec16<T>::N;
  // flat parameter layout (vector_field.py:56-61)
  static constexpr int OFF_W1 = 0, OFF_B1 = OFF_W1 + H1 * NIN, OFF_W2 = OFF_B1 + H1, OFF_B2 = OFF_W2 + H2 * H1,
                       OFF_W3 = OFF_B2 + H2, OFF_B3 = OFF_W3 + NOUT * H2, NPARAM = OFF_B3 + NOUT;
  // packed shared-memory rows: [W row | bias | 0-pad], layer 1: [W1x row | w1t | b1 | 0-pad]
  static constexpr int S1 = round_up(D + 2, VN), S2 = round_up(H1 + 1, VN), S3 = round_up(H2 + 1, VN);
  static constexpr int SM_W1 = 0, SM_W2 = SM_W1 + H1 * S1, SM_W3 = SM_W2 + H2 * S2, SM_WEND = SM_W3 + NOUT * S3;
  typedef TileCfg<H1, D + 2> TC1;   // C1[k][c]: c<D space cols, c=D time col, c=D+1 bias
  typedef TileCfg<H2, H1 + 1> TC2;  // C2[k][m], m=H1 bias
  typedef TileCfg<NOUT, H2 + 1> TC3;
  static constexpr int LMAX = TC1::MP > TC2::MP ? (TC1::MP > TC3::MP ? TC1::MP : TC3::MP) : (TC2::MP > TC3::MP ? TC2::MP : TC3::MP);
  static constexpr int RMAX = TC1::NCP > TC2::NCP ? (TC1::NCP > TC3::NCP ? TC1::NCP : TC3::NCP) : (TC2::NCP > TC3::NCP ? TC2::NCP : TC3::NCP);
  static constexpr int LSTR = odd16_stride<T>(LMAX), RSTR = odd16_stride<T>(RMAX);
  static constexpr int ACC1 = 0, ACC2 = ACC1 + TC1::ACC, ACC3 = ACC2 + TC2::ACC, ACC_WARP = ACC3 + TC3::ACC;
  static constexpr int WARP_SMEM = 32 * LSTR + 32 * RSTR + ACC_WARP;  // elements of T per warp (train modes)
  static constexpr int SM_TIMES = round_up(SM_WEND, VN);               // t_i table starts here (n_steps entries)

  static size_t smem_bytes(int mode, int n_steps) {
    size_t el = SM_TIMES + round_up(n_steps + 1, VN) + round_up(NPARAM, VN) /* raw params staging */;
    if (mode != MODE_FWD) el += (size_t)NWARP * WARP_SMEM;
    return el * sizeof(T) + 64;
  }

  template <typename S>
  struct EvalState {
    S h1[H1];
    S h2[H2];
    S U2[H2][D];
    S f[D][NB];
    S J[D][NB][D];
    S Q[D];
  };

  // ---------------- velocity forward (A.1 + A.2) ----------------
  template <typename S>
  GPF_DEV static void eval_forward(const T* __restrict__ w, const S (&x)[D], T t, T bnd, S (&v)[D], EvalState<S>& st) {
    S T1[H1][D];
#pragma unroll
    for (int k = 0; k < H1; ++k) {
      T row[S1];
      lds_row<T, S1>(w + SM_W1 + k * S1, row);
      S a = S(row[D + 1] + row[D] * t);
#pragma unroll
      for (int j = 0; j < D; ++j) a = a + row[j] * x[j];
      S h = act_tanh(a);
      st.h1[k] = h;
      S sp = T(1) - h * h;
#pragma unroll
      for (int j = 0; j < D; ++j) T1[k][j] = sp * row[j];
    }
    GPF_V1_ROLL_PRAGMA
    for (int k = 0; k < H2; ++k) {
      T row[S2];
      lds_row<T, S2>(w + SM_W2 + k * S2, row);
      S a = S(row[H1]);
      S u[D];
#pragma unroll
      for (int j = 0; j < D; ++j) u[j] = S(T(0));
#pragma unroll
      for (int m = 0; m < H1; ++m) {
        a = a + row[m] * st.h1[m];
#pragma unroll
        for (int j = 0; j < D; ++j) u[j] = u[j] + row[m] * T1[m][j];
      }
      st.h2[k] = act_tanh(a);
#pragma unroll
      for (int j = 0; j < D; ++j) st.U2[k][j] = u[j];
    }
    S T2[H2][D];
#pragma unroll
    for (int m = 0; m < H2; ++m) {
      S sp = T(1) - st.h2[m] * st.h2[m];
#pragma unroll
      for (int j = 0; j < D; ++j) T2[m][j] = sp * st.U2[m][j];
    }
#pragma unroll
    for (int i = 0; i < D; ++i) {
#pragma unroll
      for (int n = 0; n < NB; ++n) {
        if (n <= i) {
          T row[S3];
          lds_row<T, S3>(w + SM_W3 + (i * NB + n) * S3, row);
          S fv = S(row[H2]);
          S Jv[D];
#pragma unroll
          for (int j = 0; j < D; ++j) Jv[j] = S(T(0));
#pragma unroll
          for (int m = 0; m < H2; ++m) {
            fv = fv + row[m] * st.h2[m];
#pragma unroll
            for (int j = 0; j < D; ++j)
              if (j >= n) Jv[j] = Jv[j] + row[m] * T2[m][j];
          }
          st.f[i][n] = fv;
#pragma unroll
          for (int j = 0; j < D; ++j) st.J[i][n][j] = Jv[j];
        }
      }
    }
    S s[D];
#pragma unroll
    for (int j = 0; j < D; ++j) s[j] = x[j] * x[j] - bnd;
#pragma unroll
    for (int i = 0; i < D; ++i) {
      S Q = S(T(0));
#pragma unroll
      for (int n = 0; n < NB; ++n) {
        if (n <= i) {
#pragma unroll
          for (int j = 0; j < D; ++j) {
            if (j >= n && j != i)
              Q = Q + (T(2) * (st.f[i][n] - st.f[j][n])) * x[j] + s[j] * (st.J[i][n][j] - st.J[j][n][j]);
          }
        }
      }
      st.Q[i] = Q;
      v[i] = s[i] * Q;
    }
  }

  // velocity / vjp_at are the units of code reuse: NOT inlined (GPF_V1_NOINLINE), so that the Euler kernels hold one
  // copy of the dual-number forward and forward+reverse evaluation instead of 10 + 4 (53k SASS instructions, 850 KB:
  // ncu showed 14 no-instruction stall cycles per issued instruction).  Arguments are D-vectors: the call costs a few
  // local-memory moves.
  template <typename S>
  GPF_V1_CALL static void velocity(const T* __restrict__ w, const S (&x)[D], T t, T bnd, S (&v)[D]) {
    EvalState<S> st;
    eval_forward<S>(w, x, t, bnd, v, st);
  }

  // ---------------- warp-level weight-gradient sink ----------------
  struct Sink {
    T* sL;    // [32][LSTR]
    T* sR;    // [32][RSTR]
    T* accs;  // lane-private accumulators of this warp
    int lane;
  };

  template <class TC>
  GPF_DEV static void gemm_accum(const Sink& sk, T (&acc)[TC::TPL][16]) {
    if (TC::NT >= 32) {
#pragma unroll
      for (int tp = 0; tp < TC::TPL; ++tp) {
        const int tile = sk.lane + 32 * tp;
        if (tile < TC::NT) {
          const int ti = tile / TC::TN, tj = tile % TC::TN;
#pragma unroll 4
          for (int s = 0; s < 32; ++s) {
            T l[4], r[4];
            lds_row<T, 4>(sk.sL + s * LSTR + 4 * ti, l);
            lds_row<T, 4>(sk.sR + s * RSTR + 4 * tj, r);
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
              for (int b = 0; b < 4; ++b) acc[tp][a * 4 + b] += l[a] * r[b];
          }
        }
      }
    } else {
      if (sk.lane < TC::NT * TC::KS) {
        const int tile = sk.lane % TC::NT, slice = sk.lane / TC::NT;
        const int ti = tile / TC::TN, tj = tile % TC::TN;
#pragma unroll 4
        for (int q = 0; q < 32 / TC::KS; ++q) {
          const int s = slice + q * TC::KS;
          T l[4], r[4];
          lds_row<T, 4>(sk.sL + s * LSTR + 4 * ti, l);
          lds_row<T, 4>(sk.sR + s * RSTR + 4 * tj, r);
#pragma unroll
          for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[0][a * 4 + b] += l[a] * r[b];
        }
      }
    }
  }

  // one rank-32 update term: every lane contributes the row pair (L, R) of its sample
  template <class TC, int LC, int RC, typename S>
  GPF_DEV static void sink_term(const Sink& sk, const S (&L)[TC::M], const S (&R)[TC::NC], T (&acc)[TC::TPL][16]) {
    T lrow[TC::MP], rrow[TC::NCP];
#pragma unroll
    for (int i = 0; i < TC::MP; ++i) lrow[i] = i < TC::M ? comp<LC>(L[i < TC::M ? i : 0]) : T(0);
#pragma unroll
    for (int i = 0; i < TC::NCP; ++i) rrow[i] = i < TC::NC ? comp<RC>(R[i < TC::NC ? i : 0]) : T(0);
    sts_row<T, TC::MP>(sk.sL + sk.lane * LSTR, lrow);
    sts_row<T, TC::NCP>(sk.sR + sk.lane * RSTR, rrow);
    __syncwarp();
    gemm_accum<TC>(sk, acc);
    __syncwarp();
  }
  // real scalars: one term; dual scalars: epsilon part of the product = L.e R.v + L.v R.e
  template <class TC, typename S>
  GPF_DEV static void sink_pair(const Sink& sk, const S (&L)[TC::M], const S (&R)[TC::NC], T (&acc)[TC::TPL][16]) {
    if (IsDual<S>::value) {
      sink_term<TC, 1, 0, S>(sk, L, R, acc);
      sink_term<TC, 0, 1, S>(sk, L, R, acc);
    } else {
      sink_term<TC, 0, 0, S>(sk, L, R, acc);
    }
  }
  template <class TC>
  GPF_DEV static void acc_zero(T (&acc)[TC::TPL][16]) {
#pragma unroll
    for (int tp = 0; tp < TC::TPL; ++tp)
#pragma unroll
      for (int e = 0; e < 16; ++e) acc[tp][e] = T(0);
  }
  template <class TC>
  GPF_DEV static void acc_flush(const Sink& sk, int acc_off, const T (&acc)[TC::TPL][16]) {
    // layout: chunk index (tp*CH + c)*32 + lane, chunk = one 16-byte vector
    constexpr int CH = 16 / VN;
#pragma unroll
    for (int tp = 0; tp < TC::TPL; ++tp)
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        T* p = sk.accs + acc_off + ((tp * CH + c) * 32 + sk.lane) * VN;
        T cur[VN];
        lds_row<T, VN>(p, cur);
#pragma unroll
        for (int q = 0; q < VN; ++q) cur[q] += acc[tp][c * VN + q];
        sts_row<T, VN>(p, cur);
      }
  }
  // read back entry (r, c) of layer accumulator summed over the K-slices of one warp
  template <class TC>
  GPF_DEV static T acc_read(const T* accs_warp, int acc_off, int r, int c) {
    constexpr int CH = 16 / VN;
    const int tile = (r / 4) * TC::TN + (c / 4);
    const int e = (r % 4) * 4 + (c % 4);
    T sum = T(0);
    if (TC::NT >= 32) {
      const int lane = tile % 32, tp = tile / 32;
      sum = accs_warp[acc_off + ((tp * CH + e / VN) * 32 + lane) * VN + e % VN];
    } else {
#pragma unroll
      for (int q = 0; q < TC::KS; ++q) {
        const int lane = tile + TC::NT * q;
        sum += accs_warp[acc_off + ((e / VN) * 32 + lane) * VN + e % VN];
      }
    }
    return sum;
  }

  // ---------------- velocity reverse sweep (A.3 + A.4) ----------------
  // Given the state of eval_forward at (x, t) and the cotangent vbar of v, returns xbar = J_v^T vbar and
  // feeds the parameter-gradient contractions to the sink.
  template <typename S>
  GPF_DEV static void eval_reverse(const T* __restrict__ w, const Sink& sk, const S (&x)[D], T t, T bnd,
                                   const EvalState<S>& st, const S (&vbar)[D], S (&xbar)[D]) {
    S s[D], sb[D], Qb[D];
#pragma unroll
    for (int j = 0; j < D; ++j) {
      s[j] = x[j] * x[j] - bnd;
      Qb[j] = vbar[j] * s[j];
      sb[j] = vbar[j] * st.Q[j];
      xbar[j] = S(T(0));
    }
    S fb[NOUT];
    S Jb[D][NOUT];  // Jb[j][p]
#pragma unroll
    for (int p = 0; p < NOUT; ++p) {
      fb[p] = S(T(0));
#pragma unroll
      for (int j = 0; j < D; ++j) Jb[j][p] = S(T(0));
    }
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
      for (int n = 0; n < NB; ++n)
        if (n <= i) {
#pragma unroll
          for (int j = 0; j < D; ++j)
            if (j >= n && j != i) {
              S c2 = (T(2) * x[j]) * Qb[i];
              fb[i * NB + n] += c2;
              fb[j * NB + n] -= c2;
              xbar[j] += (T(2) * (st.f[i][n] - st.f[j][n])) * Qb[i];
              sb[j] += (st.J[i][n][j] - st.J[j][n][j]) * Qb[i];
              S c3 = s[j] * Qb[i];
              Jb[j][i * NB + n] += c3;
              Jb[j][j * NB + n] -= c3;
            }
        }
#pragma unroll
    for (int j = 0; j < D; ++j) xbar[j] += (T(2) * x[j]) * sb[j];

    // ---- output layer
    S sp2[H2];
#pragma unroll
    for (int m = 0; m < H2; ++m) sp2[m] = T(1) - st.h2[m] * st.h2[m];
    {
      T acc[TC3::TPL][16];
      acc_zero<TC3>(acc);
      S R[H2 + 1];
#pragma unroll
      for (int m = 0; m < H2; ++m) R[m] = st.h2[m];
      R[H2] = S(T(1));
      sink_pair<TC3, S>(sk, fb, R, acc);
#pragma unroll
      for (int j = 0; j < D; ++j) {
#pragma unroll
        for (int m = 0; m < H2; ++m) R[m] = sp2[m] * st.U2[m][j];
        R[H2] = S(T(0));
        sink_pair<TC3, S>(sk, Jb[j], R, acc);
      }
      acc_flush<TC3>(sk, ACC3, acc);
    }
    S hb2[H2], Tb2[H2][D];
#pragma unroll
    for (int m = 0; m < H2; ++m) {
      hb2[m] = S(T(0));
#pragma unroll
      for (int j = 0; j < D; ++j) Tb2[m][j] = S(T(0));
    }
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
      for (int n = 0; n < NB; ++n)
        if (n <= i) {
          const int p = i * NB + n;
          T row[S3];
          lds_row<T, S3>(w + SM_W3 + p * S3, row);
#pragma unroll
          for (int m = 0; m < H2; ++m) {
            hb2[m] += row[m] * fb[p];
#pragma unroll
            for (int j = 0; j < D; ++j)
              if (j >= n) Tb2[m][j] += row[m] * Jb[j][p];
          }
        }
    // ---- hidden layer 2
    S ab2[H2], Ub2[D][H2];
#pragma unroll
    for (int k = 0; k < H2; ++k) {
      S rs = S(T(0));
#pragma unroll
      for (int j = 0; j < D; ++j) rs += Tb2[k][j] * st.U2[k][j];
      S spp = (T(-2) * st.h2[k]) * sp2[k];
      ab2[k] = hb2[k] * sp2[k] + rs * spp;
#pragma unroll
      for (int j = 0; j < D; ++j) Ub2[j][k] = sp2[k] * Tb2[k][j];
    }
    S sp1[H1];
#pragma unroll
    for (int m = 0; m < H1; ++m) sp1[m] = T(1) - st.h1[m] * st.h1[m];
    {
      T acc[TC2::TPL][16];
      acc_zero<TC2>(acc);
      S R[H1 + 1];
#pragma unroll
      for (int m = 0; m < H1; ++m) R[m] = st.h1[m];
      R[H1] = S(T(1));
      sink_pair<TC2, S>(sk, ab2, R, acc);
#pragma unroll
      for (int j = 0; j < D; ++j) {
#pragma unroll
        for (int m = 0; m < H1; ++m) R[m] = sp1[m] * w[SM_W1 + m * S1 + j];
        R[H1] = S(T(0));
        sink_pair<TC2, S>(sk, Ub2[j], R, acc);
      }
      acc_flush<TC2>(sk, ACC2, acc);
    }
    S hb1[H1], Tb1[H1][D];
#pragma unroll
    for (int m = 0; m < H1; ++m) {
      hb1[m] = S(T(0));
#pragma unroll
      for (int j = 0; j < D; ++j) Tb1[m][j] = S(T(0));
    }
    GPF_V1_ROLL_PRAGMA
    for (int k = 0; k < H2; ++k) {
      T row[S2];
      lds_row<T, S2>(w + SM_W2 + k * S2, row);
#pragma unroll
      for (int m = 0; m < H1; ++m) {
        hb1[m] += row[m] * ab2[k];
#pragma unroll
        for (int j = 0; j < D; ++j) Tb1[m][j] += row[m] * Ub2[j][k];
      }
    }
    // ---- hidden layer 1
    S ab1[H1], Ub1[D][H1];
#pragma unroll
    for (int k = 0; k < H1; ++k) {
      T row[S1];
      lds_row<T, S1>(w + SM_W1 + k * S1, row);
      S rs = S(T(0));
#pragma unroll
      for (int j = 0; j < D; ++j) rs += Tb1[k][j] * row[j];
      S spp = (T(-2) * st.h1[k]) * sp1[k];
      ab1[k] = hb1[k] * sp1[k] + rs * spp;
#pragma unroll
      for (int j = 0; j < D; ++j) {
        Ub1[j][k] = sp1[k] * Tb1[k][j];
        xbar[j] += row[j] * ab1[k];
      }
    }
    {
      T acc[TC1::TPL][16];
      acc_zero<TC1>(acc);
      S R[D + 2];
#pragma unroll
      for (int j = 0; j < D; ++j) R[j] = x[j];
      R[D] = S(t);
      R[D + 1] = S(T(1));
      sink_pair<TC1, S>(sk, ab1, R, acc);
#pragma unroll
      for (int j = 0; j < D; ++j) {
#pragma unroll
        for (int c = 0; c < D + 2; ++c) R[c] = S(c == j ? T(1) : T(0));
        // R is a constant unit vector: only the L.e x R.v term is non-zero for duals
        if (IsDual<S>::value) sink_term<TC1, 1, 0, S>(sk, Ub1[j], R, acc);
        else sink_term<TC1, 0, 0, S>(sk, Ub1[j], R, acc);
      }
      acc_flush<TC1>(sk, ACC1, acc);
    }
  }

  // J_v(x,t)^T vbar with recomputed activations
  template <typename S>
  GPF_V1_CALL static void vjp_at(const T* __restrict__ w, const Sink& sk, const S (&x)[D], T t, T bnd, const S (&vbar)[D],
                             S (&xbar)[D]) {
    EvalState<S> st;
    S v[D];
    eval_forward<S>(w, x, t, bnd, v, st);
    eval_reverse<S>(w, sk, x, t, bnd, st, vbar, xbar);
  }

  // ---------------- Euler penalization (A.7) ----------------
  // epsilon part of a(x) = (v(x + delta v(x,t), t+delta) - v(x,t)) / delta along direction `dir`
  GPF_DEV static void pen_tangent(const T* __restrict__ w, const T (&x)[D], T t, T bnd, T delta, const T (&dir)[D],
                                  T (&ae)[D]) {
    typedef Dual<T> DS;
    DS X[D], v1[D], X2[D], v2[D];
#pragma unroll
    for (int j = 0; j < D; ++j) X[j] = DS(x[j], dir[j]);
    velocity<DS>(w, X, t, bnd, v1);
#pragma unroll
    for (int j = 0; j < D; ++j) X2[j] = X[j] + delta * v1[j];
    velocity<DS>(w, X2, t + delta, bnd, v2);
#pragma unroll
    for (int j = 0; j < D; ++j) ae[j] = (v2[j].e - v1[j].e) / delta;
  }
  // reverse of pen_tangent for one direction: ubar = cotangent of the epsilon part of a.  Accumulates
  // d/dx into xbar and the parameter gradients into the sink.
  GPF_DEV static void pen_reverse(const T* __restrict__ w, const Sink& sk, const T (&x)[D], T t, T bnd, T delta,
                                  const T (&dir)[D], const T (&ubar)[D], T (&xbar)[D]) {
    typedef Dual<T> DS;
    DS X[D], v1[D], X2[D];
#pragma unroll
    for (int j = 0; j < D; ++j) X[j] = DS(x[j], dir[j]);
    velocity<DS>(w, X, t, bnd, v1);
#pragma unroll
    for (int j = 0; j < D; ++j) X2[j] = X[j] + delta * v1[j];
    // dual adjoints: (.v) = cotangent of the tangent part, (.e) = cotangent of the value part
    DS V2b[D], X2b[D], V1b[D], X1b[D];
#pragma unroll
    for (int j = 0; j < D; ++j) V2b[j] = DS(ubar[j] / delta, T(0));
    vjp_at<DS>(w, sk, X2, t + delta, bnd, V2b, X2b);
#pragma unroll
    for (int j = 0; j < D; ++j) V1b[j] = DS(-ubar[j] / delta, T(0)) + delta * X2b[j];
    vjp_at<DS>(w, sk, X, t, bnd, V1b, X1b);
#pragma unroll
    for (int j = 0; j < D; ++j) xbar[j] += X2b[j].e + X1b[j].e;
  }
};

// =====================================================================================================
// the kernel
// =====================================================================================================
template <typename T, int D, int NB, int H1, int H2, int BLOCK, int MODE, bool EULER>
#ifndef GPF_SMALL_MINB
#define GPF_SMALL_MINB 1
#endif
__global__ void __launch_bounds__(BLOCK, GPF_SMALL_MINB) gpflow_small_kernel(const KArgs<T> a) {
  typedef SmallKernel<T, D, NB, H1, H2, BLOCK> K;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint64_t bar;
  T* const smem = reinterpret_cast<T*>(smem_raw);
  T* const w = smem;                                   // packed weights
  T* const times = smem + K::SM_TIMES;                 // t_i, i = 0..n_steps
  T* const raw = times + round_up(a.n_steps + 1, K::VN);  // flat params as staged by TMA
  T* const warp_area = raw + round_up(K::NPARAM, K::VN);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // ---- prologue: TMA bulk copy of the flat parameters, re-pack into padded rows, time table ----
  stage_params_to_smem<T>(raw, a.params, K::NPARAM, &bar);
  for (int i = tid; i < K::SM_WEND; i += BLOCK) {
    T val = T(0);
    if (i < K::SM_W2) {
      const int k = i / K::S1, c = i % K::S1;
      if (c < K::NIN) val = raw[K::OFF_W1 + k * K::NIN + c];
      else if (c == K::NIN) val = raw[K::OFF_B1 + k];
    } else if (i < K::SM_W3) {
      const int k = (i - K::SM_W2) / K::S2, c = (i - K::SM_W2) % K::S2;
      if (c < H1) val = raw[K::OFF_W2 + k * H1 + c];
      else if (c == H1) val = raw[K::OFF_B2 + k];
    } else {
      const int k = (i - K::SM_W3) / K::S3, c = (i - K::SM_W3) % K::S3;
      if (c < H2) val = raw[K::OFF_W3 + k * H2 + c];
      else if (c == H2) val = raw[K::OFF_B3 + k];
    }
    w[i] = val;
  }
  if (tid == 0) {
    T t = a.t0;  // repeated addition in the working precision (ode_solver.py:80, res_net.py:262)
    for (int i = 0; i <= a.n_steps; ++i) {
      times[i] = t;
      t = t + a.dt;
    }
  }
  typename K::Sink sk;
  sk.lane = lane;
  sk.sL = sk.sR = sk.accs = nullptr;
  if (MODE != MODE_FWD) {
    T* wa = warp_area + (size_t)warp * K::WARP_SMEM;
    sk.sL = wa;
    sk.sR = wa + 32 * K::LSTR;
    sk.accs = sk.sR + 32 * K::RSTR;
    for (int i = lane; i < K::ACC_WARP; i += 32) sk.accs[i] = T(0);
  }
  __syncthreads();

  const bool wrap = (a.flags & 0x1u) != 0;
  const bool want_lagr = (a.flags & 0x2u) != 0;
  const T bnd = T(1) - Eps<T>::value;                    // velocity_field.py:96
  const T dt = a.dt, hdt = a.dt / T(2), c6 = a.dt / T(6);
  const T delta = T(2.0 * sqrt((double)Eps<T>::value));  // res_net.py:346
  const T rsq2s = T(1) / (T(1.4142135623730951) * a.sigma);
  const T sq2s = T(1.4142135623730951) * a.sigma;
  T reg_cot = T(1);
  if (MODE == MODE_BWD && a.reg_cotangent != nullptr) reg_cot = *a.reg_cotangent;
  const T pen_w = a.euler_coef * a.loss_scale;

  const long long TS = (long long)gridDim.x * BLOCK;
  const long long gtid = (long long)blockIdx.x * BLOCK + tid;
  T ot_acc = T(0), pen_acc = T(0), lagr_acc = T(0);

  for (long long base = (long long)blockIdx.x * BLOCK; base < a.B; base += TS) {
    const long long b = base + tid;
    const bool valid = b < a.B;
    T xin[D], x[D];
#pragma unroll
    for (int j = 0; j < D; ++j) {
      xin[j] = valid ? a.x_in[b * D + j] : T(0);
      x[j] = wrap ? erf_real(xin[j] * rsq2s) : xin[j];
    }
    // ---------------- forward time loop ----------------
    for (int i = 0; i < a.n_steps; ++i) {
      const T t = times[i];
      if (MODE != MODE_FWD) {
#pragma unroll
        for (int j = 0; j < D; ++j) a.traj_ws[((long long)i * D + j) * TS + gtid] = x[j];
      }
      if (a.traj_out != nullptr && valid) {
#pragma unroll
        for (int j = 0; j < D; ++j) a.traj_out[((long long)i * a.B + b) * D + j] = x[j];
      }
      // classical RK4 (ode_order_method.py:66-71) as a rolled stage loop (one inlined copy of the velocity):
      // ksum = k1 + 2 k2 + 2 k3 + k4 accumulated left to right like the reference expression
      T ys[D], ksum[D];
#pragma unroll
      for (int j = 0; j < D; ++j) { ys[j] = x[j]; ksum[j] = T(0); }
#pragma unroll 1
      for (int sg = 0; sg < 4; ++sg) {
        const T ts = sg == 0 ? t : (sg == 3 ? t + dt : t + hdt);
        T kv[D];
        K::template velocity<T>(w, ys, ts, bnd, kv);
        const T wk = (sg == 1 || sg == 2) ? T(2) : T(1);
        const T ak = sg == 2 ? dt : hdt;
#pragma unroll
        for (int j = 0; j < D; ++j) {
          ksum[j] = sg == 0 ? kv[j] : ksum[j] + wk * kv[j];
          ys[j] = x[j] + ak * kv[j];
        }
      }
#pragma unroll
      for (int j = 0; j < D; ++j) x[j] = x[j] + c6 * ksum[j];
      if (want_lagr) {  // pairs x_{i+1} with t_i (res_net.py:259-262)
        T vv[D];
        K::template velocity<T>(w, x, t, bnd, vv);
        T sq = T(0);
#pragma unroll
        for (int j = 0; j < D; ++j) sq += vv[j] * vv[j];
        if (valid) lagr_acc += T(0.5) * sq;
      }
      if (EULER && MODE == MODE_FWD) {
        T zv[D], wv[D], aw[D], az[D];
#pragma unroll
        for (int j = 0; j < D; ++j) {
          zv[j] = a.zw[(i * 2 + 0) * D + j];
          wv[j] = a.zw[(i * 2 + 1) * D + j];
        }
        K::pen_tangent(w, x, t, bnd, delta, wv, aw);
        K::pen_tangent(w, x, t, bnd, delta, zv, az);
        T q = T(0);
#pragma unroll
        for (int j = 0; j < D; ++j) q += zv[j] * aw[j] - wv[j] * az[j];
        if (valid) pen_acc += q * q;
      }
    }
    if (a.traj_out != nullptr && valid) {
#pragma unroll
      for (int j = 0; j < D; ++j) a.traj_out[((long long)a.n_steps * a.B + b) * D + j] = x[j];
    }
    // ---------------- wrapper epilogue (network.py:178-185) ----------------
    T yv[D];
    bool inband[D];
#pragma unroll
    for (int j = 0; j < D; ++j) {
      if (wrap) {
        const T lim = T(1) - Eps<T>::value;
        T gc = x[j] < -lim ? -lim : (x[j] > lim ? lim : x[j]);
        inband[j] = (x[j] >= -lim) && (x[j] <= lim);
        yv[j] = sq2s * erfinv_real(gc);
      } else {
        yv[j] = x[j];
        inband[j] = true;
      }
    }
    if (valid) {
      if (a.y_out != nullptr) {
#pragma unroll
        for (int j = 0; j < D; ++j) a.y_out[b * D + j] = yv[j];
      }
      if (a.logdet_out != nullptr) {
        T ny = T(0), nx = T(0);
#pragma unroll
        for (int j = 0; j < D; ++j) {
          ny += yv[j] * yv[j];
          nx += xin[j] * xin[j];
        }
        a.logdet_out[b] = (ny - nx) / (T(2) * a.sigma * a.sigma);
      }
    }
    if (MODE == MODE_FWD) continue;

    // ---------------- seed of the reverse sweep ----------------
    T xb[D];
#pragma unroll
    for (int j = 0; j < D; ++j) {
      T yb;
      if (MODE == MODE_TRAIN) {
        const T diff = valid ? (a.x_data[b * D + j] - yv[j]) : T(0);
        ot_acc += diff * diff;
        yb = T(-2) * diff * a.loss_scale;                      // d/dy of loss_scale * sum (x - y)^2  (network.py:82)
      } else {
        yb = valid ? a.grad_y[b * D + j] : T(0);
      }
      if (wrap) {
        // dy/dg = sqrt(pi/2) sigma exp(y^2 / (2 sigma^2)) inside the clip band, 0 outside
        const T dydg = T(1.2533141373155001) * a.sigma * exp_real(yv[j] * yv[j] / (T(2) * a.sigma * a.sigma));
        xb[j] = inband[j] ? yb * dydg : T(0);
      } else {
        xb[j] = yb;
      }
    }
    // ---------------- reverse time loop ----------------
    for (int i = a.n_steps - 1; i >= 0; --i) {
      const T t = times[i];
      if (EULER) {
        // penalty is evaluated at (x_{i+1}, t_i); x currently holds x_{i+1}
        T zv[D], wv[D], aw[D], az[D];
#pragma unroll
        for (int j = 0; j < D; ++j) {
          zv[j] = a.zw[(i * 2 + 0) * D + j];
          wv[j] = a.zw[(i * 2 + 1) * D + j];
        }
        K::pen_tangent(w, x, t, bnd, delta, wv, aw);
        K::pen_tangent(w, x, t, bnd, delta, zv, az);
        T q = T(0);
#pragma unroll
        for (int j = 0; j < D; ++j) q += zv[j] * aw[j] - wv[j] * az[j];
        if (!valid) q = T(0);
        pen_acc += q * q;
        const T g = T(2) * q * pen_w * reg_cot;  // d(pen_w q^2)/dq
        T ub[D];
#pragma unroll
        for (int j = 0; j < D; ++j) ub[j] = g * zv[j];
        K::pen_reverse(w, sk, x, t, bnd, delta, wv, ub, xb);
#pragma unroll
        for (int j = 0; j < D; ++j) ub[j] = -g * wv[j];
        K::pen_reverse(w, sk, x, t, bnd, delta, zv, ub, xb);
      }
      T x0[D];
#pragma unroll
      for (int j = 0; j < D; ++j) x0[j] = a.traj_ws[((long long)i * D + j) * TS + gtid];
      // recompute the stage points y_1..y_4 (A.5); rolled loop, one inlined copy of the velocity
      T ystage[4][D];
#pragma unroll
      for (int j = 0; j < D; ++j) ystage[0][j] = x0[j];
#pragma unroll 1
      for (int sg = 0; sg < 3; ++sg) {
        const T ts = sg == 0 ? t : t + hdt;
        T yv2[D], kv[D];
#pragma unroll
        for (int j = 0; j < D; ++j) yv2[j] = sg == 0 ? ystage[0][j] : (sg == 1 ? ystage[1][j] : ystage[2][j]);
        K::template velocity<T>(w, yv2, ts, bnd, kv);
        const T ak = sg == 2 ? dt : hdt;
#pragma unroll
        for (int j = 0; j < D; ++j) {
          const T yn = x0[j] + ak * kv[j];
          if (sg == 0) ystage[1][j] = yn;
          else if (sg == 1) ystage[2][j] = yn;
          else ystage[3][j] = yn;
        }
      }
      // adjoint of the step: kb_4 = dt/6 xb; kb_3 = dt/3 xb + dt yb_4; kb_2 = dt/3 xb + dt/2 yb_3; kb_1 = dt/6 xb + dt/2 yb_2
      T ybsum[D], ybprev[D];
#pragma unroll
      for (int j = 0; j < D; ++j) { ybsum[j] = T(0); ybprev[j] = T(0); }
#pragma unroll 1
      for (int sg = 3; sg >= 0; --sg) {
        const T ts = sg == 0 ? t : (sg == 3 ? t + dt : t + hdt);
        const T wk = (sg == 1 || sg == 2) ? T(2) * c6 : c6;
        const T ak = sg == 3 ? T(0) : (sg == 2 ? dt : hdt);
        T yv2[D], kb[D], yb[D];
#pragma unroll
        for (int j = 0; j < D; ++j) {
          yv2[j] = sg == 0 ? ystage[0][j] : (sg == 1 ? ystage[1][j] : (sg == 2 ? ystage[2][j] : ystage[3][j]));
          kb[j] = wk * xb[j] + ak * ybprev[j];
        }
        K::template vjp_at<T>(w, sk, yv2, ts, bnd, kb, yb);
#pragma unroll
        for (int j = 0; j < D; ++j) { ybprev[j] = yb[j]; ybsum[j] += yb[j]; }
      }
#pragma unroll
      for (int j = 0; j < D; ++j) {
        xb[j] = xb[j] + ybsum[j];
        x[j] = x0[j];
      }
    }
    if (a.grad_x_out != nullptr && valid) {
#pragma unroll
      for (int j = 0; j < D; ++j) {
        T g = xb[j];
        if (wrap)  // dg0/dx = sqrt(2/pi)/sigma exp(-x^2/(2 sigma^2))
          g *= T(0.7978845608028654) / a.sigma * exp_real(-xin[j] * xin[j] / (T(2) * a.sigma * a.sigma));
        a.grad_x_out[b * D + j] = g;
      }
    }
  }

  // ---------------- per-CTA partials: [ot, reg, lagr | grads] ----------------
  ot_acc = warp_sum(ot_acc);
  pen_acc = warp_sum(pen_acc);
  lagr_acc = warp_sum(lagr_acc);
  __shared__ T red[3][32];
  if (lane == 0) {
    red[0][warp] = ot_acc;
    red[1][warp] = pen_acc;
    red[2][warp] = lagr_acc;
  }
  __syncthreads();
  T* part = a.partials + (size_t)blockIdx.x * a.npart;
  if (tid < 3) {
    T s = T(0);
    for (int q = 0; q < K::NWARP; ++q) s += red[tid][q];
    if (tid == 0) s *= a.loss_scale;
    if (tid == 1) s *= pen_w;
    if (tid == 2) s *= a.dt_constraint * a.loss_scale;
    part[tid] = s;
  }
  if (MODE != MODE_FWD) {
    for (int e = tid; e < K::NPARAM; e += BLOCK) {
      int layer, r, c;
      if (e < K::OFF_B1) { layer = 1; r = e / K::NIN; c = e % K::NIN; }
      else if (e < K::OFF_W2) { layer = 1; r = e - K::OFF_B1; c = K::NIN; }
      else if (e < K::OFF_B2) { layer = 2; r = (e - K::OFF_W2) / H1; c = (e - K::OFF_W2) % H1; }
      else if (e < K::OFF_W3) { layer = 2; r = e - K::OFF_B2; c = H1; }
      else if (e < K::OFF_B3) { layer = 3; r = (e - K::OFF_W3) / H2; c = (e - K::OFF_W3) % H2; }
      else { layer = 3; r = e - K::OFF_B3; c = H2; }
      T s = T(0);
      for (int q = 0; q < K::NWARP; ++q) {
        const T* aw = warp_area + (size_t)q * K::WARP_SMEM + 32 * K::LSTR + 32 * K::RSTR;
        if (layer == 1) s += K::template acc_read<typename K::TC1>(aw, K::ACC1, r, c);
        else if (layer == 2) s += K::template acc_read<typename K::TC2>(aw, K::ACC2, r, c);
        else s += K::template acc_read<typename K::TC3>(aw, K::ACC3, r, c);
      }
      part[3 + e] = s;
    }
  }
}

// fixed-order reduction of the per-CTA partial vectors -> [scalars | grads]
template <typename T>
__global__ void gpflow_finalize_kernel(const T* __restrict__ partials, int n_cta, int npart, int count,
                                       T* __restrict__ scalars_out, T* __restrict__ grads_out) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= count) return;
  T s = T(0);
  for (int c = 0; c < n_cta; ++c) s += partials[(size_t)c * npart + e];
  if (e < 3) {
    if (scalars_out != nullptr) scalars_out[e] = s;
  } else if (grads_out != nullptr) {
    grads_out[e - 3] = s;
  }
}

// The same reduction for MANY partial vectors (sf2r / sf2e: one per resident warp, 1 776 of them): a block of 8 warps owns
// 32 consecutive entries; warp w adds the vectors c = w, w + 8, ... (coalesced 128-byte reads, independent loads), then
// the 8 partial sums are added in warp order.  Still a fixed order -> bit-reproducible; 222 dependent adds per thread
// instead of 1 776 (the serial version cost ~0.4 ms per step, 3 % of a 2^21-sample launch).
template <typename T>
__global__ void __launch_bounds__(256) gpflow_finalize_wide_kernel(const T* __restrict__ partials, int n_cta, int npart, int count,
                                                                   T* __restrict__ scalars_out, T* __restrict__ grads_out) {
  __shared__ T red[8][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int e = blockIdx.x * 32 + lane;
  T s0 = T(0), s1 = T(0), s2 = T(0), s3 = T(0);
  if (e < count) {
    int c = w;
    for (; c + 24 < n_cta; c += 32) {
      s0 += partials[(size_t)c * npart + e];
      s1 += partials[(size_t)(c + 8) * npart + e];
      s2 += partials[(size_t)(c + 16) * npart + e];
      s3 += partials[(size_t)(c + 24) * npart + e];
    }
    for (; c < n_cta; c += 8) s0 += partials[(size_t)c * npart + e];
  }
  red[w][lane] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  if (w == 0 && e < count) {
    T s = red[0][lane];
#pragma unroll
    for (int i = 1; i < 8; ++i) s += red[i][lane];
    if (e < 3) {
      if (scalars_out != nullptr) scalars_out[e] = s;
    } else if (grads_out != nullptr) {
      grads_out[e - 3] = s;
    }
  }
}

}  // namespace gpflow
