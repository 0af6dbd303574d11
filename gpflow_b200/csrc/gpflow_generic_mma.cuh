// gpflow_b200 -- tensor-core contractions of the generic (any d, nb, depth, width) fp32 kernels.
//
// The three contractions of one linear layer, per sample (one warp = one sample, rows = value row + d tangent rows):
//   forward   Y[r][k]  = sum_m X[r][m]  W[k][m]        vector_field.py:75-128   (A.1)
//   backprop  Xb[r][m] = sum_k Yb[r][k] W[k][m]                                  (A.4)
//   gradient  dW[k][m] += sum_r Yb[r][k] X[r][m]                                 (A.4)
// are issued as mma.sync.m16n8k8 TF32 tiles with the 3xTF32 error-compensated split (x = hi + lo, hi = tf32(x),
// lo = tf32(x - hi); x*y ~= lo*hi' + hi*lo' + hi*hi', fp32 accumulation): single-pass TF32 (2^-11) would break the
// 1e-5 parity bar, the dropped lo*lo' term is 2^-22 relative.  Measured on B200 (tools/mma_tf32_bench.cu): 279 TFLOP/s
// for mma.sync tf32 = 3.9x the FP32 FFMA roof, i.e. 1.3x FFMA-equivalent after the 3x split, with ~25x fewer issued
// instructions than the scalar loops these routines replace.  Operands are read straight from the warp's scratch
// rows / the nn.Linear weights with bounds predicates (zero padding of ragged tiles), so no layout change is needed.
// Fragment layouts (PTX ISA, m16n8k8 .tf32): g = lane >> 2, t = lane & 3;
//   A (16x8, row): a0 (g, t) a1 (g+8, t) a2 (g, t+4) a3 (g+8, t+4);  B (8x8, col): b0 (t, g) b1 (t+4, g);
//   C (16x8): c0 (g, 2t) c1 (g, 2t+1) c2 (g+8, 2t) c3 (g+8, 2t+1).
#pragma once
#include "gpflow_common.cuh"

#ifndef GPF_GENERIC_MMA
#define GPF_GENERIC_MMA 1  // 0: scalar FFMA contractions everywhere (the fp64 / Dual path)
#endif

namespace gpflow {

GPF_DEV void tf32_split(float x, uint32_t& hi, uint32_t& lo) {
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(x));
  const float r = x - __uint_as_float(hi);
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lo) : "f"(r));
}
GPF_DEV void mma_tf32(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
// c += a * b with the 3xTF32 split, small terms first
GPF_DEV void mma_3xtf32(float (&c)[4], const uint32_t (&ah)[4], const uint32_t (&al)[4], const uint32_t (&bh)[2],
                        const uint32_t (&bl)[2]) {
  mma_tf32(c, al, bh);
  mma_tf32(c, ah, bl);
  mma_tf32(c, ah, bh);
}

// row r of the (1 + d)-row operand: r == 0 -> value vector v0, 1 <= r <= d -> tangent rows V[(r-1) * ld ..], else zero
GPF_DEV float row_elem(const float* __restrict__ v0, const float* __restrict__ V, int ld, int d, int r, int col, int ncol) {
  if (col >= ncol || r > d) return 0.0f;
  return r == 0 ? v0[col] : V[(r - 1) * ld + col];
}

// Fire-and-forget accumulation into the warp's OWN gradient partial (no other warp touches it, and every address is
// always updated by the same lane, so the sum order -- hence the result -- stays run-to-run deterministic): a vector
// `red` instead of load + add + store takes the partial's L2 / HBM round trip off the warp's dependency chain.
// GPF_GENERIC_RED=0 restores the read-modify-write.
#ifndef GPF_GENERIC_RED
#define GPF_GENERIC_RED 1
#endif
GPF_DEV void red_add_f32(float* p, float a) {
  asm volatile("red.global.add.f32 [%0], %1;" ::"l"(__cvta_generic_to_global(p)), "f"(a) : "memory");
}
GPF_DEV void red_add_v2(float* p, float a, float b) {
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(__cvta_generic_to_global(p)), "f"(a), "f"(b) : "memory");
}
GPF_DEV void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(__cvta_generic_to_global(p)), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

constexpr int GM_NT = 8;  // n-tiles per accumulator block (64 output columns)

// Y[r][n] = sum_k X[r][k] * Bm(k, n), r < 1 + d, n < N, k < K.  TRANS_B = false: Bm(k, n) = W[n * ldw + k] (forward:
// W is [out][in], n = out, k = in); TRANS_B = true: Bm(k, n) = W[k * ldw + n] (backprop: n = in, k = out).
// Results: row 0 -> y0[n] (+ bias0[n] if bias0), rows 1..d -> Y[(r-1) * ldy + n].  One warp; ends with __syncwarp().
template <bool TRANS_B>
GPF_DEV void gm_rows_times_w(const float* __restrict__ x0, const float* __restrict__ X, int ldx, int d, int K, int N,
                             const float* __restrict__ W, int ldw, const float* __restrict__ bias0,
                             float* __restrict__ y0, float* __restrict__ Y, int ldy, int lane) {
  const int g = lane >> 2, t = lane & 3;
  const int R = d + 1;
  for (int r0 = 0; r0 < R; r0 += 16) {
    const int rl = r0 + g, rh = r0 + g + 8;
    for (int nb0 = 0; nb0 < N; nb0 += 8 * GM_NT) {
      float c[GM_NT][4];
#pragma unroll
      for (int i = 0; i < GM_NT; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.0f;
      // operands of k-step k0 are fetched one step ahead (register double buffer): the loads of step k0 + 8 are in
      // flight while the 3 x GM_NT mma of step k0 issue
      float an[4], wn[GM_NT][2];
      auto fetch = [&](int k0) {
        an[0] = row_elem(x0, X, ldx, d, rl, k0 + t, K);
        an[1] = row_elem(x0, X, ldx, d, rh, k0 + t, K);
        an[2] = row_elem(x0, X, ldx, d, rl, k0 + t + 4, K);
        an[3] = row_elem(x0, X, ldx, d, rh, k0 + t + 4, K);
#pragma unroll
        for (int i = 0; i < GM_NT; ++i) {
          const int n = nb0 + 8 * i + g;
          wn[i][0] = wn[i][1] = 0.0f;
          if (n < N) {
            if (k0 + t < K) wn[i][0] = TRANS_B ? W[(k0 + t) * ldw + n] : W[n * ldw + k0 + t];
            if (k0 + t + 4 < K) wn[i][1] = TRANS_B ? W[(k0 + t + 4) * ldw + n] : W[n * ldw + k0 + t + 4];
          }
        }
      };
      fetch(0);
      for (int k0 = 0; k0 < K; k0 += 8) {
        uint32_t ah[4], al[4], bh[GM_NT][2], bl[GM_NT][2];
#pragma unroll
        for (int q = 0; q < 4; ++q) tf32_split(an[q], ah[q], al[q]);
#pragma unroll
        for (int i = 0; i < GM_NT; ++i) {
          tf32_split(wn[i][0], bh[i][0], bl[i][0]);
          tf32_split(wn[i][1], bh[i][1], bl[i][1]);
        }
        if (k0 + 8 < K) fetch(k0 + 8);
#pragma unroll
        for (int i = 0; i < GM_NT; ++i)
          if (nb0 + 8 * i < N) mma_3xtf32(c[i], ah, al, bh[i], bl[i]);  // warp-uniform
      }
#pragma unroll
      for (int i = 0; i < GM_NT; ++i) {
        const int n = nb0 + 8 * i + 2 * t;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int r = (q & 2) ? rh : rl, nn = n + (q & 1);
          if (nn < N && r <= d) {
            if (r == 0) y0[nn] = c[i][q] + (bias0 != nullptr ? bias0[nn] : 0.0f);
            else Y[(r - 1) * ldy + nn] = c[i][q];
          }
        }
      }
    }
  }
  __syncwarp();
}

// gw[k * N + m] += sum_{r <= d} Yb[r][k] * X[r][m], k < M (outputs), m < N (inputs); rows as in row_elem.
// KS > 0: the 1 + d rows fit in KS k-steps of 8 (d <= 8 KS - 1): the A fragments of a 16-output strip (already split into
// tf32 hi / lo) are fetched once and kept in registers across all the m-tiles of the strip.  KS = 0: any d.
template <int KS>
GPF_DEV void gm_weight_grad_impl(const float* __restrict__ yb0, const float* __restrict__ Yb, int ldyb,
                                 const float* __restrict__ x0, const float* __restrict__ X, int ldx, int d, int M, int N,
                                 float* __restrict__ gw, int lane) {
  const int g = lane >> 2, t = lane & 3;
  const int R = d + 1;
  const bool full_m = (N & 7) == 0;
  for (int k0 = 0; k0 < M; k0 += 16) {
    uint32_t Ah[KS > 0 ? KS : 1][4], Al[KS > 0 ? KS : 1][4];
    if (KS > 0) {
#pragma unroll
      for (int s = 0; s < KS; ++s) {
        const int r0 = 8 * s;
        tf32_split(row_elem(yb0, Yb, ldyb, d, r0 + t, k0 + g, M), Ah[s][0], Al[s][0]);
        tf32_split(row_elem(yb0, Yb, ldyb, d, r0 + t, k0 + g + 8, M), Ah[s][1], Al[s][1]);
        tf32_split(row_elem(yb0, Yb, ldyb, d, r0 + t + 4, k0 + g, M), Ah[s][2], Al[s][2]);
        tf32_split(row_elem(yb0, Yb, ldyb, d, r0 + t + 4, k0 + g + 8, M), Ah[s][3], Al[s][3]);
      }
    }
    const bool full_k = k0 + 16 <= M;
    for (int m0 = 0; m0 < N; m0 += 8) {
      float c[4] = {0.0f, 0.0f, 0.0f, 0.0f}, go[4];
      const bool full = full_k && full_m;
#pragma unroll
      for (int q = 0; q < 4; ++q) {  // the accumulator's old values travel while the tile is computed
        const int k = k0 + g + ((q & 2) ? 8 : 0), m = m0 + 2 * t + (q & 1);
        go[q] = (full || (k < M && m < N)) ? gw[k * N + m] : 0.0f;
      }
      if (KS > 0) {
#pragma unroll
        for (int s = 0; s < KS; ++s) {
          const int r0 = 8 * s;
          uint32_t bh[2], bl[2];
          tf32_split(row_elem(x0, X, ldx, d, r0 + t, m0 + g, N), bh[0], bl[0]);
          tf32_split(row_elem(x0, X, ldx, d, r0 + t + 4, m0 + g, N), bh[1], bl[1]);
          mma_3xtf32(c, Ah[s], Al[s], bh, bl);
        }
      } else {
        for (int r0 = 0; r0 < R; r0 += 8) {
          uint32_t ah[4], al[4], bh[2], bl[2];
          // A[k][r] = Yb[r][k]
          tf32_split(row_elem(yb0, Yb, ldyb, d, r0 + t, k0 + g, M), ah[0], al[0]);
          tf32_split(row_elem(yb0, Yb, ldyb, d, r0 + t, k0 + g + 8, M), ah[1], al[1]);
          tf32_split(row_elem(yb0, Yb, ldyb, d, r0 + t + 4, k0 + g, M), ah[2], al[2]);
          tf32_split(row_elem(yb0, Yb, ldyb, d, r0 + t + 4, k0 + g + 8, M), ah[3], al[3]);
          // B[r][m] = X[r][m]
          tf32_split(row_elem(x0, X, ldx, d, r0 + t, m0 + g, N), bh[0], bl[0]);
          tf32_split(row_elem(x0, X, ldx, d, r0 + t + 4, m0 + g, N), bh[1], bl[1]);
          mma_3xtf32(c, ah, al, bh, bl);
        }
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int k = k0 + g + ((q & 2) ? 8 : 0), m = m0 + 2 * t + (q & 1);
        if (full || (k < M && m < N)) gw[k * N + m] = go[q] + c[q];
      }
    }
  }
}
// Aligned fast path of gm_weight_grad (M % 16 == 0, N % 64 == 0, gw 8-byte aligned): one 16-output strip x 64-input block
// per pass -- the A fragment of a k-step (4 loads, 4 splits) feeds 8 m-tiles, no column predicates, the 16 x 64 accumulator
// block is read-modify-written once with 64-bit accesses.  ~13 instead of ~35 (+ predicates) instructions per 3 mma.
GPF_DEV void gm_weight_grad_fast(const float* __restrict__ yb0, const float* __restrict__ Yb, int ldyb,
                                 const float* __restrict__ x0, const float* __restrict__ X, int ldx, int d, int M, int N,
                                 float* __restrict__ gw, int lane, bool red) {
  const int g = lane >> 2, t = lane & 3;
  const int R = d + 1;
  for (int k0 = 0; k0 < M; k0 += 16) {
    for (int mb = 0; mb < N; mb += 64) {
      float c[8][4];
#pragma unroll
      for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.0f;
      for (int r0 = 0; r0 < R; r0 += 8) {
        const int ra = r0 + t, rb = r0 + t + 4;      // rb >= 4: always a tangent row
        const bool va = ra <= d, vb = rb <= d;
        const float* __restrict__ pa = ra == 0 ? yb0 : Yb + (size_t)(ra - 1) * ldyb;
        const float* __restrict__ pb = Yb + (size_t)(rb - 1) * ldyb;
        const float* __restrict__ qa = (ra == 0 ? x0 : X + (size_t)(ra - 1) * ldx) + mb + g;
        const float* __restrict__ qb = X + (size_t)(rb - 1) * ldx + mb + g;
        float av[4], bv[8][2];
        av[0] = va ? pa[k0 + g] : 0.0f; av[1] = va ? pa[k0 + g + 8] : 0.0f;
        av[2] = vb ? pb[k0 + g] : 0.0f; av[3] = vb ? pb[k0 + g + 8] : 0.0f;
#pragma unroll
        for (int i = 0; i < 8; ++i) { bv[i][0] = va ? qa[8 * i] : 0.0f; bv[i][1] = vb ? qb[8 * i] : 0.0f; }
        uint32_t ah[4], al[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) tf32_split(av[q], ah[q], al[q]);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          uint32_t bh[2], bl[2];
          tf32_split(bv[i][0], bh[0], bl[0]);
          tf32_split(bv[i][1], bh[1], bl[1]);
          mma_3xtf32(c[i], ah, al, bh, bl);
        }
      }
      float* __restrict__ g0 = gw + (size_t)(k0 + g) * N + mb + 2 * t;
      float* __restrict__ g1 = g0 + (size_t)8 * N;
      if (GPF_GENERIC_RED && red) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          red_add_v2(g0 + 8 * i, c[i][0], c[i][1]);
          red_add_v2(g1 + 8 * i, c[i][2], c[i][3]);
        }
        continue;
      }
      float2 o0[8], o1[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        o0[i] = *reinterpret_cast<const float2*>(g0 + 8 * i);
        o1[i] = *reinterpret_cast<const float2*>(g1 + 8 * i);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        *reinterpret_cast<float2*>(g0 + 8 * i) = make_float2(o0[i].x + c[i][0], o0[i].y + c[i][1]);
        *reinterpret_cast<float2*>(g1 + 8 * i) = make_float2(o1[i].x + c[i][2], o1[i].y + c[i][3]);
      }
    }
  }
}
GPF_DEV void gm_weight_grad(const float* __restrict__ yb0, const float* __restrict__ Yb, int ldyb, const float* __restrict__ x0,
                            const float* __restrict__ X, int ldx, int d, int M, int N, float* __restrict__ gw, int lane,
                            bool red) {
  if ((M & 15) == 0 && (N & 63) == 0 && (reinterpret_cast<uintptr_t>(gw) & 7) == 0)
    gm_weight_grad_fast(yb0, Yb, ldyb, x0, X, ldx, d, M, N, gw, lane, red);
  else if (d <= 7) gm_weight_grad_impl<1>(yb0, Yb, ldyb, x0, X, ldx, d, M, N, gw, lane);
  else if (d <= 15) gm_weight_grad_impl<2>(yb0, Yb, ldyb, x0, X, ldx, d, M, N, gw, lane);
  else gm_weight_grad_impl<0>(yb0, Yb, ldyb, x0, X, ldx, d, M, N, gw, lane);
}

// ---------------------------------------------------------------------------------------------------
// fp32 last-layer passes with 128-bit accesses: lane owns input units m = 4 lane + 128 c.  Same arithmetic as the
// scalar loops of Generic<T>::last_layer_forward / last_layer_reverse (see the formulas there); 4 kept rows per pass,
// so 8-12 independent 16-byte loads are in flight per lane -- the passes are L2 / HBM latency bound.
// ---------------------------------------------------------------------------------------------------
GPF_DEV float4 gl_ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
GPF_DEV void gl_st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
GPF_DEV float gl_dot4(float4 a, float4 b, float acc) { return fmaf(a.x, b.x, fmaf(a.y, b.y, fmaf(a.z, b.z, fmaf(a.w, b.w, acc)))); }
GPF_DEV float4 gl_fma4(float s, float4 a, float4 acc) {
  return make_float4(fmaf(s, a.x, acc.x), fmaf(s, a.y, acc.y), fmaf(s, a.z, acc.z), fmaf(s, a.w, acc.w));
}

// VW-wide lane vectors for the last-layer passes: VW = 4 (128-bit accesses, 128 input units per warp pass) or VW = 2
// (64-bit, 64 units per pass -- all 32 lanes busy when H <= 64, e.g. the zero-padded C3 shape)
template <int VW> struct GLv { float x[VW]; };
template <int VW> GPF_DEV GLv<VW> gv_zero() { GLv<VW> r;
#pragma unroll
  for (int i = 0; i < VW; ++i) r.x[i] = 0.0f;
  return r; }
template <int VW> GPF_DEV GLv<VW> gv_ld(const float* p);
template <> GPF_DEV GLv<4> gv_ld<4>(const float* p) { const float4 v = *reinterpret_cast<const float4*>(p); GLv<4> r; r.x[0] = v.x; r.x[1] = v.y; r.x[2] = v.z; r.x[3] = v.w; return r; }
template <> GPF_DEV GLv<2> gv_ld<2>(const float* p) { const float2 v = *reinterpret_cast<const float2*>(p); GLv<2> r; r.x[0] = v.x; r.x[1] = v.y; return r; }
GPF_DEV void gv_st(float* p, const GLv<4>& a) { *reinterpret_cast<float4*>(p) = make_float4(a.x[0], a.x[1], a.x[2], a.x[3]); }
GPF_DEV void gv_st(float* p, const GLv<2>& a) { *reinterpret_cast<float2*>(p) = make_float2(a.x[0], a.x[1]); }
GPF_DEV void gv_red(float* p, const GLv<4>& a) { red_add_v4(p, a.x[0], a.x[1], a.x[2], a.x[3]); }
GPF_DEV void gv_red(float* p, const GLv<2>& a) { red_add_v2(p, a.x[0], a.x[1]); }
template <int VW> GPF_DEV float gv_dot(const GLv<VW>& a, const GLv<VW>& b, float acc) {   // same nesting as gl_dot4
#pragma unroll
  for (int i = VW - 1; i >= 0; --i) acc = fmaf(a.x[i], b.x[i], acc);
  return acc; }
template <int VW> GPF_DEV GLv<VW> gv_fma(float s, const GLv<VW>& a, const GLv<VW>& acc) { GLv<VW> r;
#pragma unroll
  for (int i = 0; i < VW; ++i) r.x[i] = fmaf(s, a.x[i], acc.x[i]);
  return r; }
template <int VW> GPF_DEV GLv<VW> gv_add(const GLv<VW>& a, const GLv<VW>& b) { GLv<VW> r;
#pragma unroll
  for (int i = 0; i < VW; ++i) r.x[i] = a.x[i] + b.x[i];
  return r; }

// f_p = W[p].h + b_p, g_p = W[p].R_n, D_p = W[p].T_L[i] for every kept row p = (i, n <= min(i, nb-1))
template <int VW>
GPF_DEV void gl_dots_v(const float* __restrict__ W, const float* __restrict__ bias, const float* __restrict__ hL,
                        const float* __restrict__ TL, const float* __restrict__ R, float* __restrict__ f,
                        float* __restrict__ G, float* __restrict__ D, int d, int nb, int H, int lane) {
  for (int i = 0; i < d; ++i) {
    const int nmax = i < nb - 1 ? i : nb - 1;
    for (int n0 = 0; n0 <= nmax; n0 += 4) {
      float af[4] = {0.f, 0.f, 0.f, 0.f}, ag[4] = {0.f, 0.f, 0.f, 0.f}, ad[4] = {0.f, 0.f, 0.f, 0.f};
      for (int m = VW * lane; m < H; m += 32 * VW) {
        GLv<VW> w[4], r[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int nq = n0 + q <= nmax ? n0 + q : nmax;  // clamped rows are discarded below
          w[q] = gv_ld<VW>(W + (size_t)(i * nb + nq) * H + m);
          r[q] = gv_ld<VW>(R + nq * H + m);
        }
        const GLv<VW> hm = gv_ld<VW>(hL + m), ti = gv_ld<VW>(TL + i * H + m);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          af[q] = gv_dot<VW>(w[q], hm, af[q]);
          ag[q] = gv_dot<VW>(w[q], r[q], ag[q]);
          ad[q] = gv_dot<VW>(w[q], ti, ad[q]);
        }
      }
      const float rf = warp_sum4(af, lane), rg = warp_sum4(ag, lane), rd = warp_sum4(ad, lane);
      const int q = 2 * ((lane >> 4) & 1) + ((lane >> 3) & 1);
      if ((lane & 7) == 0 && n0 + q <= nmax) {
        const int p = i * nb + n0 + q;
        f[p] = rf + bias[p];
        G[p] = rg;
        D[p] = rd;
      }
    }
  }
}

// pass A (hb_L, D part of Tb_L, dW) and pass B (Rb_n, CR_n, R part of Tb_L and of sb) of the reverse last layer
template <int VW>
GPF_DEV void gl_reverse_v(const float* __restrict__ W, float* __restrict__ gw, const float* __restrict__ hL,
                           const float* __restrict__ TL, const float* __restrict__ R, const float* __restrict__ fb,
                           const float* __restrict__ Qb, const float* __restrict__ s, const float* __restrict__ SQ,
                           float* __restrict__ hb, float* __restrict__ Tb, float* sb, int d, int nb, int H, int lane,
                           bool red) {
  const GLv<VW> z4 = gv_zero<VW>();
  for (int m0 = 0; m0 < H; m0 += 32 * VW) {
    const int m = m0 + VW * lane;
    const bool act = m < H;
    GLv<VW> hbm = z4;
    const GLv<VW> hm = act ? gv_ld<VW>(hL + m) : z4;
    for (int i = 0; i < d; ++i) {
      const int nmax = i < nb - 1 ? i : nb - 1;
      const GLv<VW> Ti = act ? gv_ld<VW>(TL + i * H + m) : z4;
      const float qbi = Qb[i], nsi = -s[i];
      GLv<VW> Ei = z4;
      if (act) {
        for (int n0 = 0; n0 <= nmax; n0 += 4) {
          GLv<VW> w[4], go[4], r[4];
          float fbp[4], dbp[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {  // every load of the pass before its stores
            const int nq = n0 + q <= nmax ? n0 + q : nmax;
            const size_t p = (size_t)(i * nb + nq);
            w[q] = gv_ld<VW>(W + p * H + m);
            go[q] = (GPF_GENERIC_RED && red) ? z4 : gv_ld<VW>(gw + p * H + m);
            r[q] = gv_ld<VW>(R + nq * H + m);
            fbp[q] = fb[p];
            dbp[q] = nsi * SQ[nq];
          }
#pragma unroll
          for (int q = 0; q < 4; ++q)
            if (n0 + q <= nmax) {
              hbm = gv_fma<VW>(fbp[q], w[q], hbm);
              Ei = gv_fma<VW>(dbp[q], w[q], Ei);
              {
                const GLv<VW> u = gv_fma<VW>(fbp[q], hm, gv_fma<VW>(qbi, r[q], gv_fma<VW>(dbp[q], Ti, go[q])));
                if (GPF_GENERIC_RED && red) gv_red(gw + (size_t)(i * nb + n0 + q) * H + m, u);
                else gv_st(gw + (size_t)(i * nb + n0 + q) * H + m, u);
              }
            }
        }
        gv_st(Tb + i * H + m, Ei);
      }
    }
    if (act) gv_st(hb + m, hbm);
    GLv<VW> CR = z4;
    for (int n = 0; n < nb; ++n) {
      GLv<VW> rb = z4;
      if (act) {
        int i = n;
        for (; i + 4 <= d; i += 4) {
          GLv<VW> w[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) w[q] = gv_ld<VW>(W + (size_t)((i + q) * nb + n) * H + m);
#pragma unroll
          for (int q = 0; q < 4; ++q) rb = gv_fma<VW>(Qb[i + q], w[q], rb);
        }
        for (; i < d; ++i) rb = gv_fma<VW>(Qb[i], gv_ld<VW>(W + (size_t)(i * nb + n) * H + m), rb);
      }
      CR = gv_add<VW>(CR, rb);
      const int jhi = (n == nb - 1) ? d - 1 : n;
      for (int j = n; j <= jhi; ++j) {
        float part = 0.0f;
        if (act) {
          part = gv_dot<VW>(CR, gv_ld<VW>(TL + j * H + m), 0.0f);
          gv_st(Tb + j * H + m, gv_fma<VW>(s[j], CR, gv_ld<VW>(Tb + j * H + m)));
        }
        part = warp_sum(part);
        if (lane == 0) sb[j] = sb[j] + part;
      }
    }
  }
}

// Aligned fast path of gm_rows_times_w (K % 8 == 0, N % 64 == 0, ldw % 4 == 0, ldy % 4 == 0, 16-byte aligned bases):
// B(k, n) = Bhi/Blo[k * ldw + n] are the weights ALREADY split into their tf32 hi / lo parts (done once per call by
// the parameter-preparation kernel instead of per sample, per m-tile and per k-step), and the 8 n-tiles of a 64-column
// block are interleaved -- tile i, local column c is global column nb0 + 8 c + i -- so that a lane's B fragments of
// all 8 tiles are 8 consecutive floats (2 LDG.128 per k-row and part) and its C fragments 16 consecutive columns per
// row (4 STG.128).  ~65 instead of ~190 issued instructions per k-step of 24 mma.
GPF_DEV void gm_rows_times_w_fast(const float* __restrict__ x0, const float* __restrict__ X, int ldx, int d, int K, int N,
                                  const float* __restrict__ Bhi, const float* __restrict__ Blo, int ldw,
                                  const float* __restrict__ bias0, float* __restrict__ y0, float* __restrict__ Y, int ldy,
                                  int lane) {
  const int g = lane >> 2, t = lane & 3;
  const int R = d + 1;
  for (int r0 = 0; r0 < R; r0 += 16) {
    const int rl = r0 + g, rh = r0 + g + 8;
    const bool vl = rl <= d, vh = rh <= d;
    const float* __restrict__ pl = rl == 0 ? x0 : X + (size_t)(rl - 1) * ldx;
    const float* __restrict__ ph = X + (size_t)(rh - 1) * ldx;   // rh >= 8: always a tangent row
    for (int nb0 = 0; nb0 < N; nb0 += 64) {
      float c[8][4];
#pragma unroll
      for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.0f;
      float an[4];
      an[0] = vl ? pl[t] : 0.0f; an[1] = vh ? ph[t] : 0.0f; an[2] = vl ? pl[t + 4] : 0.0f; an[3] = vh ? ph[t + 4] : 0.0f;
      for (int k0 = 0; k0 < K; k0 += 8) {
        const size_t o0 = (size_t)(k0 + t) * ldw + nb0 + 8 * g, o1 = o0 + 4 * (size_t)ldw;
        float h0[8], l0[8], h1[8], l1[8];
        *reinterpret_cast<float4*>(h0) = gl_ld4(Bhi + o0); *reinterpret_cast<float4*>(h0 + 4) = gl_ld4(Bhi + o0 + 4);
        *reinterpret_cast<float4*>(l0) = gl_ld4(Blo + o0); *reinterpret_cast<float4*>(l0 + 4) = gl_ld4(Blo + o0 + 4);
        *reinterpret_cast<float4*>(h1) = gl_ld4(Bhi + o1); *reinterpret_cast<float4*>(h1 + 4) = gl_ld4(Bhi + o1 + 4);
        *reinterpret_cast<float4*>(l1) = gl_ld4(Blo + o1); *reinterpret_cast<float4*>(l1 + 4) = gl_ld4(Blo + o1 + 4);
        uint32_t ah[4], al[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) tf32_split(an[q], ah[q], al[q]);
        if (k0 + 8 < K) {
          an[0] = vl ? pl[k0 + 8 + t] : 0.0f; an[1] = vh ? ph[k0 + 8 + t] : 0.0f;
          an[2] = vl ? pl[k0 + 12 + t] : 0.0f; an[3] = vh ? ph[k0 + 12 + t] : 0.0f;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t bh[2] = {__float_as_uint(h0[i]), __float_as_uint(h1[i])};
          const uint32_t bl[2] = {__float_as_uint(l0[i]), __float_as_uint(l1[i])};
          mma_3xtf32(c[i], ah, al, bh, bl);
        }
      }
      // lane's 16 consecutive columns of rows rl / rh: [c0 of tiles 0..7 | c1 of tiles 0..7]
      const int col = nb0 + 16 * t;
#pragma unroll
      for (int hrow = 0; hrow < 2; ++hrow) {
        const int r = hrow ? rh : rl;
        if (r <= d) {
          float v[16];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            v[i] = c[i][2 * hrow];
            v[8 + i] = c[i][2 * hrow + 1];
          }
          float* __restrict__ dst = r == 0 ? y0 + col : Y + (size_t)(r - 1) * ldy + col;
          if (r == 0 && bias0 != nullptr) {
#pragma unroll
            for (int q = 0; q < 16; ++q) v[q] += bias0[col + q];
          }
#pragma unroll
          for (int q = 0; q < 16; q += 4) gl_st4(dst + q, make_float4(v[q], v[q + 1], v[q + 2], v[q + 3]));
        }
      }
    }
  }
  __syncwarp();
}

}  // namespace gpflow
