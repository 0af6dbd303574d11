// gpflow_b200 -- "tc" path, part 2: the GEMM ops on the tcgen05 engine and the host-side orchestration of a call.
// See gpflow_tc.cuh for the decomposition and the layouts, gpflow_tc_gemm.cuh for the engine.
#pragma once
#include <stdlib.h>

#include "gpflow_tc.cuh"

// 1: the 64-wide layer GEMMs issue 2 instead of 3 tcgen05.mma per K step (Op::N_CAT, gpflow_tc_gemm.cuh).  Parity-green,
// measured on the engine test at M = 180 224, N = K = 64 (the C3 tangent layer): 30 us against 28 us, C3 step 60.7 against
// 60.1 ms -- the MMA thread is then idle waiting for the epilogue, which reads twice the TMEM columns; the tile time at this
// width (~3.7 kclk per 128 x 64 x 64) is set by the shared-memory path (operand stores, MMA operand reads, epilogue
// staging), not by the instruction count.  Kept as a build switch.
#ifndef GPF_TC_NCAT
#define GPF_TC_NCAT 0
#endif

namespace gpflow {
namespace tc {

// ---------------------------------------------------------------------------------------------------------------
// hidden layer GEMM: out[r][:] = epilogue(A[r][:] W^T), W (pre-split) by TMA.  Used for
//   mode 0  value rows     h_l = tanh(h_{l-1} W_l^T + b_l), sp_l = 1 - h_l^2                  (vector_field.py:84-88)
//   mode 1  tangent rows   T_l[j] = sp_l (.) (T_{l-1}[j] W_l^T), rows r = j S + s              (vector_field.py:119-126)
//   mode 2  plain          reverse through W_l: [Tb_{l-1}; hb_{l-1}] = [Ub_l; ab_l] W_l  (B operand = W_l^T copy)
//   mode 3  last layer F   F[q][s] = W_{L+1}[q] . h_L[s] + b[q] for ALL output rows q (N tiles), column-layout store
// ---------------------------------------------------------------------------------------------------------------
template <int NT_, int ST_>
struct OpLayer {
  static constexpr int NT = NT_, STAGES = ST_;
  static constexpr bool B_TMA = true;
  static constexpr bool N_CAT = NT_ <= 64 && GPF_TC_NCAT;   // narrow layers are bound by the MMA instruction count
  static constexpr int A_MODE = MODE_ROWS, B_MODE = MODE_ROWS;
  struct Params {
    CUtensorMap map;
    const float* A;
    float* out;
    float* out2;
    const float* aux;
    const float* bias;
    long long M;
    int K, N, lo_row, mode, S, ntn;   // ntn: number of N tiles (1 for the hidden layers)
    const float* gen_sp;              // non-null (layer 2 tangent rows): A is not read but generated on the fly,
    const float* gen_w;               //   A[(j, s)][k] = gen_sp[s][k] * gen_w[j][k]   (T_1[j] = sp_1 (.) W1x[:, j])
  };
  struct Tile { int nk, n, n0; long long m0; };
  GPF_DEV static int num_tiles(const Params& p) { return (int)((p.M + MT - 1) / MT) * p.ntn; }
  GPF_DEV static void decode(const Params& p, int t, Tile& tl) {
    tl.m0 = (long long)(t / p.ntn) * MT;
    tl.n0 = (t % p.ntn) * NT;
    tl.nk = p.K / KC;
    const int rem = p.N - tl.n0;
    tl.n = rem >= NT ? NT : (rem + 15) / 16 * 16;
  }
  GPF_DEV static const CUtensorMap* tmap(const Params& p, const Tile&) { return &p.map; }
  GPF_DEV static void tma_coords(const Params& p, const Tile& tl, int kc, int& c0, int& r_hi, int& r_lo) {
    c0 = kc * KC; r_hi = tl.n0; r_lo = p.lo_row + tl.n0;
  }
  GPF_DEV static void fetch_a(const Params& p, const Tile& tl, int kc, int ptid, float* frag) {
    if (p.gen_sp != nullptr) {   // a tile never straddles two tangent directions (S is a multiple of 128)
      const int j = (int)(tl.m0 / p.S), s0 = (int)(tl.m0 % p.S);
#pragma unroll
      for (int it = 0; it < MT * 8 / kGroupThreads; ++it) {
        const int q = ptid + it * kGroupThreads;
        const int r = q >> 3, c = q & 7;
        const float4 a = __ldg(reinterpret_cast<const float4*>(p.gen_sp + (size_t)(s0 + r) * p.K + kc * KC + c * 4));
        const float4 w = __ldg(reinterpret_cast<const float4*>(p.gen_w + (size_t)j * p.K + kc * KC + c * 4));
        frag[4 * it] = a.x * w.x; frag[4 * it + 1] = a.y * w.y; frag[4 * it + 2] = a.z * w.z; frag[4 * it + 3] = a.w * w.w;
      }
      return;
    }
    fetch_rows<MT>(frag, ptid, p.A, p.K, tl.m0, p.M, kc * KC, p.K - kc * KC);
  }
  GPF_DEV static void epilogue(const Params& p, const Tile& tl, const EpiCtx& ctx, int row, int c0, int nc, float (&v)[32]) {
    const long long r = tl.m0 + row;
    const bool valid = r < p.M;
    auto rowptr = [&](float* base) {
      return [&p, &tl, base](int rr) -> float* { const long long R = tl.m0 + rr; return R < p.M ? base + R * p.N : nullptr; };
    };
    if (p.mode == 3) {   // column layout: lanes = consecutive samples -> coalesced
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const int q = tl.n0 + c0 + i;
        if (i < nc && q < p.N && valid) p.out[(size_t)q * p.S + r] = v[i] + p.bias[q];
      }
      return;
    }
    if (p.mode == 0) {
      float sp[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const float h = (i < nc) ? act_tanh(v[i] + p.bias[c0 + i]) : 0.f;
        v[i] = h;
        sp[i] = 1.f - h * h;
      }
      ctx.store_rows(v, nc, c0, rowptr(p.out), false);
      ctx.store_rows(sp, nc, c0, rowptr(p.out2), false);
      return;
    }
    if (p.mode == 1 && valid) {
      const float4* a4 = reinterpret_cast<const float4*>(p.aux + (size_t)(r % p.S) * p.N + c0);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (4 * i < nc) {
          const float4 a = __ldg(a4 + i);
          v[4 * i] *= a.x; v[4 * i + 1] *= a.y; v[4 * i + 2] *= a.z; v[4 * i + 3] *= a.w;
        }
    }
    ctx.store_rows(v, nc, c0, rowptr(p.out), false);
  }
};

// ---------------------------------------------------------------------------------------------------------------
// last layer, forward, contracted form (header of Generic<T>): grouped GEMMs
//   D_(i,n) = W[(i,n)] . T_L[i]      groups i: rows n <= min(i, nb-1), contiguous in W3s
//   G_(i,n) = W[(i,n)] . R_n         groups n: rows i >= n, contiguous in W3Gs
// (F_(i,n) = W[(i,n)] . h_L + b for all rows is one dense GEMM: OpLayer mode 3.)
// Orientation: a tcgen05.mma costs ~100-130 cycles whatever its N (measured; the A tile is re-read from shared memory
// for every instruction), so the SAMPLES go on the 256-wide N side and the group's weight rows -- at most d <= 64 of
// the 128 M rows are used -- on the M side: out[q][s], A = weight rows by TMA, B = activation rows by the producers.
// Outputs in column layout [p = i nb + n][s]: an output row is contiguous.
// ---------------------------------------------------------------------------------------------------------------
template <int ST_>
struct OpLast {
  static constexpr int NT = 256, STAGES = ST_;
  static constexpr bool A_TMA = true, B_TMA = false;
  static constexpr int A_MODE = MODE_ROWS, B_MODE = MODE_ROWS;
  struct Params {
    CUtensorMap map3, mapG;     // boxes of 128 weight rows
    const float *TL, *R;        // TL: [(d+1) S][K]
    float *G, *D;
    int d, nb, S, K, nt, lo3, loG;
  };
  struct Tile { int nk, n, fam, idx, cnt, s0; };
  GPF_DEV static int num_tiles(const Params& p) { return (p.d + p.nb) * p.nt; }
  GPF_DEV static void decode(const Params& p, int t, Tile& tl) {
    const int grp = t / p.nt;
    tl.s0 = (t % p.nt) * NT;
    tl.nk = p.K / KC;
    tl.n = NT;
    if (grp < p.d) {
      tl.fam = 1; tl.idx = grp;
      tl.cnt = (tl.idx < p.nb - 1 ? tl.idx : p.nb - 1) + 1;
    } else {
      tl.fam = 2; tl.idx = grp - p.d;
      tl.cnt = p.d - tl.idx;
    }
  }
  GPF_DEV static const CUtensorMap* tmap_a(const Params& p, const Tile& tl) { return tl.fam == 2 ? &p.mapG : &p.map3; }
  GPF_DEV static void tma_coords_a(const Params& p, const Tile& tl, int kc, int& c0, int& r_hi, int& r_lo) {
    c0 = kc * KC;
    if (tl.fam == 2) { r_hi = tl.idx * p.d + tl.idx; r_lo = p.loG + r_hi; }
    else { r_hi = tl.idx * p.nb; r_lo = p.lo3 + r_hi; }
  }
  GPF_DEV static void fetch_b(const Params& p, const Tile& tl, int kc, int ptid, float* frag) {
    const float* base = tl.fam == 1 ? p.TL + (size_t)tl.idx * p.S * p.K : p.R + (size_t)tl.idx * p.S * p.K;
    fetch_rows<NT>(frag, ptid, base, p.K, tl.s0, p.S, kc * KC, p.K - kc * KC);
  }
  GPF_DEV static void epilogue(const Params& p, const Tile& tl, const EpiCtx& ctx, int, int c0, int nc, float (&v)[32]) {
    ctx.store_rows(v, nc, c0, [&p, &tl](int q) -> float* {
      if (q >= tl.cnt) return nullptr;
      return tl.fam == 1 ? p.D + (size_t)(tl.idx * p.nb + q) * p.S + tl.s0
                         : p.G + (size_t)((tl.idx + q) * p.nb + tl.idx) * p.S + tl.s0; }, false);
  }
};

// ---------------------------------------------------------------------------------------------------------------
// last layer, reverse, data side: M = samples, N = hidden units, K = output rows (small)
//   fam 0  hb (segment q)  = sum_{p in segment} fb_p W[p],       fb_(i,n) = 2 (Qb_i SX_n - x_i SQ_n) on kept rows
//   fam 1  Rb_n            = sum_{i >= n} Qb_i W[(i,n)]
//   fam 2  TbD_j           = -s_j sum_{n <= min(j, nb-1)} SQ_n W[(j,n)]       (Db_(j,n) = -s_j SQ_n)
// A is generated from the per-sample column vectors; B = transposed weight copies by TMA.
// ---------------------------------------------------------------------------------------------------------------
template <int NT_, int ST_>
struct OpLB {
  static constexpr int NT = NT_, STAGES = ST_;
  static constexpr bool B_TMA = true;
  static constexpr int A_MODE = MODE_UNITS, B_MODE = MODE_ROWS;
  struct Params {
    CUtensorMap mapT, mapGT, mapJT;
    const float *Qb, *SX, *SQ, *x, *sj;
    float *hbp, *Rb, *TbD;
    int d, nb, nout, S, N, HTp, dpad, mt, nseg, nchunk_p;
  };
  struct Tile { int nk, n, fam, idx, cnt, m0, kc0; };
  GPF_DEV static int num_tiles(const Params& p) { return (p.nseg + p.nb + p.d) * p.mt; }
  GPF_DEV static void decode(const Params& p, int t, Tile& tl) {
    const int grp = t / p.mt;
    tl.m0 = (t % p.mt) * MT;
    tl.n = p.N;
    tl.kc0 = 0; tl.cnt = 0;
    if (grp < p.nseg) {
      tl.fam = 0; tl.idx = grp; tl.kc0 = grp * TC_SEG;
      tl.nk = p.nchunk_p - tl.kc0 < TC_SEG ? p.nchunk_p - tl.kc0 : TC_SEG;
    } else if (grp < p.nseg + p.nb) {
      tl.fam = 1; tl.idx = grp - p.nseg; tl.nk = p.dpad / KC;
    } else {
      tl.fam = 2; tl.idx = grp - p.nseg - p.nb;
      tl.cnt = (tl.idx < p.nb - 1 ? tl.idx : p.nb - 1) + 1;
      tl.nk = (tl.cnt + KC - 1) / KC;
    }
  }
  GPF_DEV static const CUtensorMap* tmap(const Params& p, const Tile& tl) {
    return tl.fam == 1 ? &p.mapGT : (tl.fam == 2 ? &p.mapJT : &p.mapT);
  }
  GPF_DEV static void tma_coords(const Params& p, const Tile& tl, int kc, int& c0, int& r_hi, int& r_lo) {
    if (tl.fam == 0) { c0 = (tl.kc0 + kc) * KC; r_hi = 0; r_lo = p.HTp; }
    else if (tl.fam == 1) { c0 = kc * KC; r_hi = tl.idx * 2 * p.HTp; r_lo = r_hi + p.HTp; }
    else { c0 = kc * KC; r_hi = tl.idx * 2 * p.HTp; r_lo = r_hi + p.HTp; }
  }
  GPF_DEV static void fetch_a(const Params& p, const Tile& tl, int kc, int ptid, float* frag) {
    const int S = p.S;
    fetch_units<MT>(frag, ptid, [&](int m, int kq, float (&v)[8]) {
      const int s = tl.m0 + m;
      if (tl.fam == 0) {
        int q = (tl.kc0 + kc) * KC + kq * 8;
        int i = q / p.nb, n = q - i * p.nb;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          float val = 0.f;
          if (q < p.nout && n <= i)
            val = 2.f * (p.Qb[(size_t)i * S + s] * p.SX[(size_t)n * S + s] - p.x[(size_t)i * S + s] * p.SQ[(size_t)n * S + s]);
          v[e] = val;
          ++q;
          if (++n == p.nb) { n = 0; ++i; }
        }
      } else if (tl.fam == 1) {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const int i = kc * KC + kq * 8 + e;
          v[e] = (i >= tl.idx && i < p.d) ? p.Qb[(size_t)i * S + s] : 0.f;
        }
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const int n = kc * KC + kq * 8 + e;
          v[e] = n < tl.cnt ? p.SQ[(size_t)n * S + s] : 0.f;
        }
      }
    });
  }
  GPF_DEV static void epilogue(const Params& p, const Tile& tl, const EpiCtx& ctx, int row, int c0, int nc, float (&v)[32]) {
    float* base;
    if (tl.fam == 0) base = p.hbp + (size_t)tl.idx * p.S * p.N;
    else if (tl.fam == 1) base = p.Rb + (size_t)tl.idx * p.S * p.N;
    else {
      base = p.TbD + (size_t)tl.idx * p.S * p.N;
      const float sc = -p.sj[(size_t)tl.idx * p.S + tl.m0 + row];
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] *= sc;
    }
    ctx.store_rows(v, nc, c0, [&p, &tl, base](int rr) -> float* { return base + (size_t)(tl.m0 + rr) * p.N; }, false);
  }
};

// ---------------------------------------------------------------------------------------------------------------
// last layer, weight gradient: out[q][k] = sum_s coef_q[s] X[s][k]  (M = output rows q, N = hidden units k, K = samples)
//   F tiles (128 rows each, all rows):  dW[(i,n)] += sum_s fb_(i,n) h_L       fb = 2 (Qb_i SX_n - x_i SQ_n) on kept rows
//   D groups i:                         dW[(i,n)] += sum_s Db_(i,n) T_L[i]    Db = -s_i SQ_n
//   G groups n:                         dW[(i,n)] += sum_s Qb_i R_n
// Same orientation argument as OpLast: the 256-wide side is N.  K is cut into segments of 8 chunks (256 samples); a
// work item = (output tile, K split) belongs to ONE CTA which adds its segments in order into the split's partial with
// red.global.add (round-to-nearest fp32) => deterministic; the partials are summed by tc_finalize_kernel.
// ---------------------------------------------------------------------------------------------------------------
template <int NT_, int ST_>
struct OpW3 {
  static constexpr int NT = NT_, STAGES = ST_;
  static constexpr bool B_TMA = false;
  static constexpr int A_MODE = MODE_ROWS, B_MODE = MODE_UNITS;
  struct Params {
    const float *TL, *R;
    const float *Qb, *SX, *SQ, *x, *sj;
    float* gW3;        // [3 families][nsplit][nout][K] partial accumulators (the families write the same rows)
    int d, nb, nout, S, K, nF, n_out, nseg, nsplit, items, G, segs_per_item, rounds;
  };
  struct Tile { int nk, n, fam, idx, cnt, chunk0, sp; };
  GPF_DEV static int num_tiles(const Params& p) { return p.rounds * p.G; }
  GPF_DEV static void decode(const Params& p, int t, Tile& tl) {
    const int r = t / p.G, b = t % p.G;
    const int item = b + (r / p.segs_per_item) * p.G;
    tl.nk = 0; tl.n = p.K; tl.fam = 0; tl.idx = 0; tl.cnt = 0; tl.chunk0 = 0; tl.sp = 0;
    if (item >= p.items) return;
    const int o = item / p.nsplit;
    tl.sp = item % p.nsplit;
    const int seg = tl.sp + (r % p.segs_per_item) * p.nsplit;
    if (seg >= p.nseg) return;
    tl.chunk0 = seg * TC_SEG;
    tl.nk = TC_SEG;
    if (o < p.nF) { tl.fam = 0; tl.idx = o; tl.cnt = p.nout - o * MT < MT ? p.nout - o * MT : MT; }
    else if (o < p.nF + p.d) { tl.fam = 1; tl.idx = o - p.nF; tl.cnt = (tl.idx < p.nb - 1 ? tl.idx : p.nb - 1) + 1; }
    else { tl.fam = 2; tl.idx = o - p.nF - p.d; tl.cnt = p.d - tl.idx; }
  }
  GPF_DEV static void fetch_a(const Params& p, const Tile& tl, int kc, int ptid, float* frag) {
    const int S = p.S;
#pragma unroll
    for (int it = 0; it < MT * 8 / kGroupThreads; ++it) {
      const int idx = ptid + it * kGroupThreads;
      const int q = idx >> 3, c = idx & 7;
      const int sg = (tl.chunk0 + kc) * KC + c * 4;
      float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
      if (q < tl.cnt) {
        if (tl.fam == 0) {
          const int prow = tl.idx * MT + q, i = prow / p.nb, n = prow - i * p.nb;
          if (n <= i) {
            const float4 qb = __ldg(reinterpret_cast<const float4*>(p.Qb + (size_t)i * S + sg));
            const float4 sx = __ldg(reinterpret_cast<const float4*>(p.SX + (size_t)n * S + sg));
            const float4 xx = __ldg(reinterpret_cast<const float4*>(p.x + (size_t)i * S + sg));
            const float4 sq = __ldg(reinterpret_cast<const float4*>(p.SQ + (size_t)n * S + sg));
            o.x = 2.f * (qb.x * sx.x - xx.x * sq.x); o.y = 2.f * (qb.y * sx.y - xx.y * sq.y);
            o.z = 2.f * (qb.z * sx.z - xx.z * sq.z); o.w = 2.f * (qb.w * sx.w - xx.w * sq.w);
          }
        } else if (tl.fam == 1) {
          const float4 sv = __ldg(reinterpret_cast<const float4*>(p.sj + (size_t)tl.idx * S + sg));
          const float4 sq = __ldg(reinterpret_cast<const float4*>(p.SQ + (size_t)q * S + sg));
          o.x = -sv.x * sq.x; o.y = -sv.y * sq.y; o.z = -sv.z * sq.z; o.w = -sv.w * sq.w;
        } else {
          o = __ldg(reinterpret_cast<const float4*>(p.Qb + (size_t)(tl.idx + q) * S + sg));
        }
      }
      frag[4 * it] = o.x; frag[4 * it + 1] = o.y; frag[4 * it + 2] = o.z; frag[4 * it + 3] = o.w;
    }
  }
  GPF_DEV static void fetch_b(const Params& p, const Tile& tl, int kc, int ptid, float* frag) {
    const int S = p.S, K = p.K;
    const float* X = tl.fam == 0 ? p.TL + (size_t)p.d * S * K : (tl.fam == 1 ? p.TL + (size_t)tl.idx * S * K : p.R + (size_t)tl.idx * S * K);
    fetch_units<NT>(frag, ptid, [&](int k, int kq, float (&v)[8]) {
      const size_t s0 = (size_t)(tl.chunk0 + kc) * KC + kq * 8;
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] = k < K ? __ldg(X + (s0 + e) * K + k) : 0.f;
    });
  }
  GPF_DEV static void epilogue(const Params& p, const Tile& tl, const EpiCtx& ctx, int, int c0, int nc, float (&v)[32]) {
    float* base = p.gW3 + (size_t)(tl.fam * p.nsplit + tl.sp) * p.nout * p.K;
    ctx.store_rows(v, nc, c0, [&p, &tl, base](int q) -> float* {
      if (q >= tl.cnt) return nullptr;
      const int prow = tl.fam == 0 ? tl.idx * MT + q : (tl.fam == 1 ? tl.idx * p.nb + q : (tl.idx + q) * p.nb + tl.idx);
      return base + (size_t)prow * p.K; }, true);
  }
};

// ---------------------------------------------------------------------------------------------------------------
// The same product for a 64-wide last hidden layer (C3: d = 10, nb = 9, H = 64), on the CUDA cores.  There the GEMM above
// runs 128-row output tiles for groups of at most d rows and spends its time in the fp32 operand fetch (175 us per launch,
// 14 % of the C3 step); the arithmetic itself is 2e8 FMAs.  Here a CTA owns (16 output rows of one family / group, one
// sample split): the coefficients of 64 samples are built once in shared memory, each lane keeps two hidden units
// (one 8-byte load per sample row) and 8 output rows in registers; the 4 sample quarters of the CTA are summed in a fixed
// order and added to the split's partial plane -- rows of one family are owned by one CTA per split, so no atomics.
// ---------------------------------------------------------------------------------------------------------------
inline bool tc_l1b_small_enabled() {   // GPFLOW_TC_L1B_SMALL=0: the round-1 first-layer gradient kernel (A/B switch)
  static const bool on = [] { const char* e = getenv("GPFLOW_TC_L1B_SMALL"); return !(e && e[0] == '0'); }();
  return on;
}
inline bool tc_l1b_small_forced() {    // GPFLOW_TC_L1B_SMALL=2: also for wide first layers (experiment)
  static const bool on = [] { const char* e = getenv("GPFLOW_TC_L1B_SMALL"); return e && e[0] == '2'; }();
  return on;
}
inline bool tc_w3_small_enabled() {   // GPFLOW_TC_W3_SMALL=0: the tensor-core GEMM for every width (A/B switch)
  static const bool on = [] { const char* e = getenv("GPFLOW_TC_W3_SMALL"); return !(e && e[0] == '0'); }();
  return on;
}
constexpr int W3S_ROWS = 16, W3S_CHUNK = 64, W3S_MAXSPLIT = 64;
struct W3SmallParams {
  const float *TL, *R, *Qb, *SX, *SQ, *x, *sj;
  float* gW3;    // [3 families][nsplit][nout][64]
  int d, nb, nout, S, nF, nkept, nDt, nGt, nsplit, spb;
};
__global__ void __launch_bounds__(256) tc_w3_small_kernel(const W3SmallParams p) {
  constexpr int K = 64;
  __shared__ __align__(16) float coef[2][W3S_ROWS][W3S_CHUNK];
  __shared__ __align__(16) float red[4][W3S_ROWS][K];
  __shared__ int rmap[W3S_ROWS][3];   // tile row -> (first index, second index, output row)
  const int S = p.S, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rgrp = warp & 1, sq = warp >> 1;   // tile rows r = 2 j + rgrp: a short group loads both row groups evenly
  int o = blockIdx.x, fam, idx, row0;
  if (o < p.nF) { fam = 0; idx = 0; row0 = o * W3S_ROWS; }
  else if (o < p.nF + p.d * p.nDt) { o -= p.nF; fam = 1; idx = o / p.nDt; row0 = (o % p.nDt) * W3S_ROWS; }
  else { o -= p.nF + p.d * p.nDt; fam = 2; idx = o / p.nGt; row0 = (o % p.nGt) * W3S_ROWS; }
  // F: only the kept rows (i, n <= min(i, nb - 1)) are enumerated -- the others have a zero coefficient
  const int cnt_all = fam == 0 ? p.nkept : (fam == 1 ? (idx < p.nb - 1 ? idx : p.nb - 1) + 1 : p.d - idx);
  const int cnt = cnt_all - row0 < W3S_ROWS ? cnt_all - row0 : W3S_ROWS;
  const int s_begin = blockIdx.y * p.spb, s_end = s_begin + p.spb < S ? s_begin + p.spb : S;
  if (cnt <= 0 || s_begin >= S) return;
  const float* X = fam == 0 ? p.TL + (size_t)p.d * S * K : (fam == 1 ? p.TL + (size_t)idx * S * K : p.R + (size_t)idx * S * K);
  if (tid < cnt) {
    const int q = row0 + tid;
    int a, b, prow;
    if (fam == 0) {
      int i = 0, before = 0;
      for (;; ++i) {
        const int c = (i < p.nb - 1 ? i : p.nb - 1) + 1;
        if (q < before + c) break;
        before += c;
      }
      a = i; b = q - before; prow = i * p.nb + b;
    } else if (fam == 1) { a = idx; b = q; prow = idx * p.nb + q; }
    else { a = idx + q; b = idx; prow = (idx + q) * p.nb + idx; }
    rmap[tid][0] = a; rmap[tid][1] = b; rmap[tid][2] = prow;
  }
  __syncthreads();
  // coefficient of (row r, sample s): thread (s = tid & 63) fills rows tid >> 6, + 4, + 8, + 12
  const int cs = tid & 63, cr = tid >> 6;
  auto fill = [&](int buf, int s0) {
    const int sg = s0 + cs;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = cr + 4 * j;
      float v = 0.f;
      if (r < cnt) {
        const int a = rmap[r][0], b = rmap[r][1];
        if (fam == 0)
          v = 2.f * (__ldg(p.Qb + (size_t)a * S + sg) * __ldg(p.SX + (size_t)b * S + sg) -
                     __ldg(p.x + (size_t)a * S + sg) * __ldg(p.SQ + (size_t)b * S + sg));
        else if (fam == 1)
          v = -__ldg(p.sj + (size_t)a * S + sg) * __ldg(p.SQ + (size_t)b * S + sg);
        else
          v = __ldg(p.Qb + (size_t)a * S + sg);
      }
      coef[buf][r][cs] = v;
    }
  };
  float2 acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = make_float2(0.f, 0.f);
  fill(0, s_begin);
  __syncthreads();
  int buf = 0;
  for (int s0 = s_begin; s0 < s_end; s0 += W3S_CHUNK, buf ^= 1) {
    if (s0 + W3S_CHUNK < s_end) fill(buf ^ 1, s0 + W3S_CHUNK);
    const float* xr = X + (size_t)(s0 + sq * 16) * K + 2 * lane;
#pragma unroll
    for (int g4 = 0; g4 < 4; ++g4) {
      float2 xv[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) xv[e] = __ldg(reinterpret_cast<const float2*>(xr + (size_t)(4 * g4 + e) * K));
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (2 * j + rgrp < cnt) {   // warp-uniform
          const float4 c = *reinterpret_cast<const float4*>(&coef[buf][2 * j + rgrp][sq * 16 + 4 * g4]);
          acc[j].x = fmaf(c.x, xv[0].x, acc[j].x); acc[j].y = fmaf(c.x, xv[0].y, acc[j].y);
          acc[j].x = fmaf(c.y, xv[1].x, acc[j].x); acc[j].y = fmaf(c.y, xv[1].y, acc[j].y);
          acc[j].x = fmaf(c.z, xv[2].x, acc[j].x); acc[j].y = fmaf(c.z, xv[2].y, acc[j].y);
          acc[j].x = fmaf(c.w, xv[3].x, acc[j].x); acc[j].y = fmaf(c.w, xv[3].y, acc[j].y);
        }
    }
    __syncthreads();
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) *reinterpret_cast<float2*>(&red[sq][2 * j + rgrp][2 * lane]) = acc[j];
  __syncthreads();
  float* plane = p.gW3 + (size_t)(fam * p.nsplit + blockIdx.y) * p.nout * K;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int e = tid + 256 * j, r = e >> 6, k = e & 63;
    if (r < cnt) {
      const float v = ((red[0][r][k] + red[1][r][k]) + red[2][r][k]) + red[3][r][k];
      plane[(size_t)rmap[r][2] * K + k] += v;
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// hidden layer weight gradient: dW_l[k][m] += sum_r Ub_l[r][k] T_{l-1}[r][m], r over the (d+1) S rows of the slab
// (tangent rows and, in row block d, ab_l against h_{l-1}).  Split-K over the CTAs: CTA b = (128-row tile km, split sp)
// owns the partial gWp[b] and adds its segments to it in order.
// ---------------------------------------------------------------------------------------------------------------
template <int NT_, int ST_>
struct OpW {
  static constexpr int NT = NT_, STAGES = ST_;
  static constexpr bool B_TMA = false;
  static constexpr int A_MODE = MODE_UNITS, B_MODE = MODE_UNITS;
  struct Params {
    const float* Ub;   // [(d+1) S][Ml]
    const float* Tm;   // [(d+1) S][Nl]
    float* gWp;        // [G][128][Nl]
    int Ml, Nl, nsplit, G, cps, nchunks, nseg;
    const float* gen_sp;   // non-null (l = 2): the tangent rows of T_1 are generated, T_1[(j, s)][n] = gen_sp[s][n] * gen_w[j][n];
    const float* gen_w;    //   only row block d of Tm (h_1) is read
    int S, d;
  };
  struct Tile { int nk, n, km, b, chunk0; };
  GPF_DEV static int num_tiles(const Params& p) { return p.G * p.nseg; }
  GPF_DEV static void decode(const Params& p, int t, Tile& tl) {
    const int seg = t / p.G;
    tl.b = t % p.G;
    tl.km = tl.b / p.nsplit;
    const int sp = tl.b % p.nsplit;
    tl.chunk0 = sp * p.cps + seg * TC_SEG;
    int hi = sp * p.cps + p.cps;
    if (hi > p.nchunks) hi = p.nchunks;
    if (hi > tl.chunk0 + TC_SEG) hi = tl.chunk0 + TC_SEG;
    tl.nk = hi > tl.chunk0 ? hi - tl.chunk0 : 0;
    tl.n = p.Nl;
  }
  GPF_DEV static void fetch_a(const Params& p, const Tile& tl, int kc, int ptid, float* frag) {
    fetch_units<MT>(frag, ptid, [&](int m, int kq, float (&v)[8]) {
      const int k = tl.km * MT + m;
      const size_t r0 = (size_t)(tl.chunk0 + kc) * KC + kq * 8;
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] = k < p.Ml ? __ldg(p.Ub + (r0 + e) * p.Ml + k) : 0.f;
    });
  }
  GPF_DEV static void fetch_b(const Params& p, const Tile& tl, int kc, int ptid, float* frag) {
    fetch_units<NT>(frag, ptid, [&](int n, int kq, float (&v)[8]) {
      const size_t r0 = (size_t)(tl.chunk0 + kc) * KC + kq * 8;
      if (p.gen_sp != nullptr && r0 < (size_t)p.d * p.S) {   // 8 consecutive rows share their tangent direction (S % 8 == 0)
        const int j = (int)(r0 / p.S), s0 = (int)(r0 % p.S);
        const float w = n < p.Nl ? __ldg(p.gen_w + (size_t)j * p.Nl + n) : 0.f;
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = n < p.Nl ? w * __ldg(p.gen_sp + (size_t)(s0 + e) * p.Nl + n) : 0.f;
        return;
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] = n < p.Nl ? __ldg(p.Tm + (r0 + e) * p.Nl + n) : 0.f;
    });
  }
  GPF_DEV static void epilogue(const Params& p, const Tile& tl, const EpiCtx& ctx, int, int c0, int nc, float (&v)[32]) {
    float* base = p.gWp + (size_t)tl.b * MT * p.Nl;
    ctx.store_rows(v, nc, c0, [&p, base](int rr) -> float* { return base + (size_t)rr * p.Nl; }, true);
  }
};

// gather the padded gradient accumulators into the caller's flat layout; scalars
struct TcTabs { const float* gWp[TC_MAXL + 1]; const float* gb[TC_MAXL + 1]; int ns[TC_MAXL + 1]; int ns3; };
__global__ void tc_finalize_kernel(TcDims g, const float* __restrict__ gW1, TcTabs tabs,
                                   const float* __restrict__ gW3, const float* __restrict__ gb3, const float* __restrict__ scal,
                                   float loss_scale, float lagr_scale, float* __restrict__ grads_out, float* __restrict__ scalars_out) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < 3 && scalars_out != nullptr)
    scalars_out[e] = e == 0 ? scal[0] * loss_scale : (e == 1 ? 0.f : scal[2] * lagr_scale);
  if (e >= g.P || grads_out == nullptr) return;
  float val = 0.f;
  for (int li = 0; li <= g.L; ++li) {
    const int ni = g.H[li], no = g.H[li + 1];
    if (e >= g.woff[li] && e < g.woff[li] + ni * no) {
      const int q = e - g.woff[li], k = q / ni, m = q % ni;
      if (li == 0) val = gW1[k * (g.d + 2) + m];
      else if (li == g.L) { for (int sp = 0; sp < tabs.ns3; ++sp) val += gW3[((size_t)sp * g.nout + k) * g.Hp[g.L] + m]; }
      else {
        const int l = li + 1, ns = tabs.ns[l], Nl = g.Hp[l - 1];
        const float* gp = tabs.gWp[l];
        const int km = k / MT, kr = k % MT;
        for (int sp = 0; sp < ns; ++sp) val += gp[((size_t)(km * ns + sp) * MT + kr) * Nl + m];
      }
      break;
    }
    if (e >= g.boff[li] && e < g.boff[li] + no) {
      const int k = e - g.boff[li];
      if (li == 0) val = gW1[k * (g.d + 2) + g.d + 1];
      else if (li == g.L) val = gb3[k];
      else val = tabs.gb[li + 1][k];
      break;
    }
  }
  grads_out[e] = val;
}

// ---------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------
struct TcPlan {
  TcDims g;
  int nslab_S;
  size_t total;
};
struct TcBuf {   // offsets in floats
  size_t W1T, bp[TC_MAXL + 1], Ws[TC_MAXL + 1], WTs[TC_MAXL + 1], W3s, W3Gs, W3Ts, W3GTs, W3JTs, weights_end;
  size_t gW1, zW1, zb[TC_MAXL + 1], zb3, gWp[TC_MAXL + 1], gb[TC_MAXL + 1], gW3, gb3, scal, grads_end;
  size_t stash, sj, sp[TC_MAXL + 1], T[TC_MAXL + 1], Ub[TC_MAXL + 1], R, Rb, hbp, F, G, D, SX, FX, SD, Q, v, ksum;
  size_t vbar, Qb, SQ, SXb, ybar, ybsum, ybprev, xb, ot_s, lg_s, total;
  int nseg_hb, nchunk_p, npad[TC_MAXL + 1], npadT[TC_MAXL + 1], wsplit_G[TC_MAXL + 1], w3_nsplit, nz, s_per_z;
  bool w3_small, l1b_small;
  int nz1;
};

inline size_t tc_per_sample_floats(const TcDims& g, int n_steps, bool train) {
  size_t f = (size_t)(4 * n_steps + 1) * g.d;
  for (int l = 1; l <= g.L; ++l) {
    f += (size_t)(g.d + 1) * g.Hp[l] + g.Hp[l];
    if (train && !(l < g.L && g.Hp[l + 1] >= g.Hp[l])) f += (size_t)(g.d + 1) * g.Hp[l];   // Ub_l (else aliased to T_{l+1})
  }
  const int nchunk_p = (g.nout + KC - 1) / KC, nseg_hb = (nchunk_p + TC_SEG - 1) / TC_SEG;
  f += (size_t)g.nb * g.Hp[g.L] * (train ? 2 : 1) + (train ? (size_t)nseg_hb * g.Hp[g.L] : 0);
  f += (size_t)3 * g.nout + 3 * g.nb + 3 * g.d + g.d;
  if (train) f += (size_t)7 * g.d + 2 * g.nb;
  return f + 2;
}
// slab size: a multiple of 256 samples, the per-slab buffers bounded by a budget, equal slabs.  The budget is sized for
// 180 GB of HBM: 7 GB lets C5 (395 KB per sample) run its 2^14-sample batch as ONE slab.  Measured at C5, 2^14 samples
// (GPFLOW_TC_SLAB_MB = 880 / 1760 / 3520 / 7040 -> slabs of 2048 / 4096 / 8192 / 16384): 1000 / 906 / 856 / 834 ms per step --
// larger slabs mean 8 x fewer launches and value-row / reduction launches that fill the chip (a 2048-row GEMM is 16 tiles).
inline int tc_slab_size(const GpFlowDesc* d, int64_t B, int op) {
  const TcDims g = tc_make_dims(d, 256);
  const size_t per = tc_per_sample_floats(g, d->n_steps, op != GPFLOW_OP_FORWARD) * 4;
  size_t budget_mb = 7040;
  if (const char* e = getenv("GPFLOW_TC_SLAB_MB")) {   // experiment knob: per-slab buffer budget
    const long long v = atoll(e);
    if (v >= 64 && v <= 65536) budget_mb = (size_t)v;
  }
  long long smax = (long long)((budget_mb << 20) / per) / 256 * 256;
  if (smax < 256) smax = 256;
  if (smax > 32768) smax = 32768;
  if (const char* e = getenv("GPFLOW_TC_SLAB")) {   // testing aid: force small slabs (multi-slab accumulation paths)
    const long long v = atoll(e) / 256 * 256;
    if (v >= 256 && v < smax) smax = v;
  }
  const long long bp = (B + 255) / 256 * 256;
  if (bp <= smax) return (int)bp;
  // equal slabs: smallest slab count that fits
  const long long nslab = (bp + smax - 1) / smax;
  long long s = ((B + nslab - 1) / nslab + 255) / 256 * 256;
  return (int)s;
}

inline void tc_carve(const TcDims& g, int n_steps, bool train, int sm_count, TcBuf& b) {
  memset(&b, 0, sizeof(b));
  size_t off = 0;
  auto take = [&](size_t n) { size_t o = off; off += (n + 255) / 256 * 256; return o; };   // 1 KB granularity
  const int S = g.S, d = g.d, L = g.L, HL = g.Hp[L];
  b.nchunk_p = (g.nout + KC - 1) / KC;
  b.nseg_hb = (b.nchunk_p + TC_SEG - 1) / TC_SEG;
  b.W1T = take((size_t)(d + 1) * g.Hp[1]);
  for (int l = 1; l <= L; ++l) b.bp[l] = take(g.Hp[l]);
  for (int l = 2; l <= L; ++l) {
    b.npad[l] = tc_round(g.Hp[l], g.HT);
    b.npadT[l] = tc_round(g.Hp[l - 1], g.HT);
    b.Ws[l] = take((size_t)2 * b.npad[l] * g.Hp[l - 1]);
    b.WTs[l] = take((size_t)2 * b.npadT[l] * g.Hp[l]);
  }
  b.W3s = take((size_t)2 * g.rows3 * HL);
  b.W3Gs = take((size_t)2 * g.rowsG * HL);
  b.W3Ts = take((size_t)2 * g.HTp * g.ldT);
  b.W3GTs = take((size_t)g.nb * 2 * g.HTp * g.dpad);
  b.W3JTs = take((size_t)g.d * 2 * g.HTp * g.nbpad);
  b.weights_end = off;
  // two-stage reductions over the samples (fixed order): nz splits of the slab, per-split planes [nz][n] that the kernels
  // add to during the whole call, summed once at the end (tc_reduce_add_kernel)
  static const int nz_max = [] { const char* e = getenv("GPFLOW_TC_NZ"); const int v = e ? atoi(e) : 0; return v >= 1 ? v : 128; }();   // C3 row: 70.8 ms at 32, 67.9 at 64, 67.6 at 128, 67.4 at 256
  b.nz = S / 64 < nz_max ? S / 64 : nz_max;
  b.s_per_z = (S + b.nz - 1) / b.nz;
  if (train) {
    b.gW1 = take((size_t)g.Hp[1] * (d + 2));
    // tc_layer1_bwd_b_small_kernel (one plane per 32 samples): narrow first layers only -- measured at C5 (d = 64, 256 wide,
    // 4 x 4 direction / unit tiles per sample block): 1056 ms per step against 993 ms with the kernel above
    b.l1b_small = ((g.Hp[1] == 64 && d <= 16) || tc_l1b_small_forced()) && g.Hp[1] % 64 == 0 && tc_l1b_small_enabled();
    b.nz1 = b.l1b_small ? S / TC_L1B_CH : b.nz;
    b.zW1 = take((size_t)b.nz1 * g.Hp[1] * (d + 2));
    b.zb3 = take((size_t)(S / 64) * g.nout);
    for (int l = 2; l <= L; ++l) b.zb[l] = take((size_t)b.nz * g.Hp[l]);
    for (int l = 2; l <= L; ++l) {
      const int mt = (g.Hp[l] + MT - 1) / MT;
      b.wsplit_G[l] = (sm_count / mt < 1 ? 1 : sm_count / mt) * mt;   // upper bound on CTAs = partial buffers
      b.gWp[l] = take((size_t)b.wsplit_G[l] * MT * g.Hp[l - 1]);
      b.gb[l] = take(g.Hp[l]);
    }
    {   // last-layer weight gradient: (output tile, K split) work items, one partial per split
      const int n_out = (g.nout + MT - 1) / MT + d + g.nb;
      const int nseg = S / (KC * TC_SEG);
      b.w3_nsplit = 2 * n_out > sm_count ? 4 : sm_count / n_out;   // many tiles: 4 splits even out the rounds
      if (b.w3_nsplit > nseg) b.w3_nsplit = nseg;
      if (b.w3_nsplit < 1) b.w3_nsplit = 1;
      b.w3_small = tc_w3_small_enabled() && HL == 64;
      if (b.w3_small) b.w3_nsplit = S / W3S_CHUNK < W3S_MAXSPLIT ? S / W3S_CHUNK : W3S_MAXSPLIT;
    }
    b.gW3 = take((size_t)3 * b.w3_nsplit * g.nout * HL);
    b.gb3 = take(g.nout);
  }
  b.scal = take(4);
  b.grads_end = off;
  b.stash = take((size_t)(4 * n_steps + 1) * d * S);
  b.sj = take((size_t)d * S);
  for (int l = 1; l <= L; ++l) {
    b.sp[l] = take((size_t)S * g.Hp[l]);
    b.T[l] = take((size_t)(d + 1) * S * g.Hp[l]);
  }
  if (train)   // Tb_l / Ub_l: T_{l+1} is dead once tb(l+1) has run, so Ub_l takes its place when it fits
    for (int l = 1; l <= L; ++l)
      b.Ub[l] = (l < L && g.Hp[l + 1] >= g.Hp[l]) ? b.T[l + 1] : take((size_t)(d + 1) * S * g.Hp[l]);
  b.R = take((size_t)g.nb * S * HL);
  if (train) {
    b.Rb = take((size_t)g.nb * S * HL);
    b.hbp = take((size_t)b.nseg_hb * S * HL);
  }
  b.F = take((size_t)g.nout * S); b.G = take((size_t)g.nout * S); b.D = take((size_t)g.nout * S);
  b.SX = take((size_t)g.nb * S); b.FX = take((size_t)g.nb * S); b.SD = take((size_t)g.nb * S);
  b.Q = take((size_t)d * S); b.v = take((size_t)d * S); b.ksum = take((size_t)d * S);
  if (train) {
    b.vbar = take((size_t)d * S); b.Qb = take((size_t)d * S); b.SQ = take((size_t)g.nb * S); b.SXb = take((size_t)g.nb * S);
    b.ybar = take((size_t)d * S); b.ybsum = take((size_t)d * S); b.ybprev = take((size_t)d * S); b.xb = take((size_t)d * S);
  }
  b.ot_s = take(S); b.lg_s = take(S);
  b.total = off;
}

inline size_t tc_workspace_bytes(const GpFlowDesc* d, int64_t B, int op) {
  const int S = tc_slab_size(d, B, op);
  const TcDims g = tc_make_dims(d, S);
  TcBuf b;
  tc_carve(g, d->n_steps, op != GPFLOW_OP_FORWARD, kMaxSMs, b);
  return b.total * sizeof(float) + 1024;
}

// GPFLOW_TC_SYNC=1 (debugging aid): synchronise after every launch and report the first failing one on stderr
inline bool tc_dbg_sync_enabled() {
  static const bool v = [] { const char* e = getenv("GPFLOW_TC_SYNC"); return e != nullptr && e[0] == '1'; }();
  return v;
}
inline bool tc_dbg_sync(cudaStream_t st, int line) {
  if (!tc_dbg_sync_enabled()) return true;
  const cudaError_t e = cudaStreamSynchronize(st);
  if (e == cudaSuccess) return true;
  fprintf(stderr, "gpflow tc: launch before gpflow_tc_ops.cuh:%d failed: %s\n", line, cudaGetErrorString(e));
  return false;
}
// launch accounting for benchmarks (gpflow_launches_total): every kernel this path enqueues
inline unsigned long long& tc_launch_counter() { static unsigned long long c = 0; return c; }
#define TC_CK(x) do { ++tc_launch_counter(); if ((x) != cudaSuccess) { cudaGetLastError(); return GPFLOW_ERR_CUDA; } if (!tc_dbg_sync(st, __LINE__)) return GPFLOW_ERR_CUDA; } while (0)
#define TC_KCK() do { ++tc_launch_counter(); if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA; if (!tc_dbg_sync(st, __LINE__)) return GPFLOW_ERR_CUDA; } while (0)

template <int HT, int ST>
int tc_run(const GpFlowDesc* d, int op, const float* params, const float* x_in, const float* x_data, const float* grad_y,
           float* y_out, float* logdet_out, float* traj_out, float* grad_x_out, float* grads_out, float* scalars_out, int64_t B,
           void* ws, size_t wsb, cudaStream_t st, int sm_count, float* v_out, double t_eval) {
  if (ws == nullptr || wsb < tc_workspace_bytes(d, B, op)) return GPFLOW_ERR_WORKSPACE;
  const bool train = op != GPFLOW_OP_FORWARD;
  const int S = tc_slab_size(d, B, op);
  const TcDims g = tc_make_dims(d, S);
  TcBuf b;
  tc_carve(g, d->n_steps, train, sm_count, b);
  float* base = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(ws) + 1023) & ~(uintptr_t)1023);
  auto P = [&](size_t o) { return base + o; };
  const int dd = g.d, L = g.L, HL = g.Hp[L], N = d->n_steps, nb = g.nb;
  const bool wrap = (d->flags & GPFLOW_FLAG_WRAP_GAUSS) != 0, want_lagr = (d->flags & GPFLOW_FLAG_LAGRANGIAN) != 0 && v_out == nullptr;
  const bool rev = (d->flags & GPFLOW_FLAG_REVERSE) != 0;
  const float dt = rev ? -(float)d->dt : (float)d->dt;
  const float t0 = rev ? (float)(d->dt * (double)d->n_steps_nominal) : 0.f;
  const float hdt = dt / 2.f, c6 = dt / 6.f;
  const float bnd = 1.f - Eps<float>::value;
  const float sigma = (float)d->sigma, sq2s = 1.4142135623730951f * sigma;

  // ---- weights: zero-fill, copies, tensor maps ----
  TC_CK(cudaMemsetAsync(base, 0, b.grads_end * sizeof(float), st));
  --tc_launch_counter();   // a memset, not a kernel
  TcWeights w;
  memset(&w, 0, sizeof(w));
  w.W1T = P(b.W1T);
  for (int l = 1; l <= L; ++l) w.bp[l] = P(b.bp[l]);
  for (int l = 2; l <= L; ++l) { w.Ws[l] = P(b.Ws[l]); w.WTs[l] = P(b.WTs[l]); w.npad[l] = b.npad[l]; w.npadT[l] = b.npadT[l]; }
  w.W3s = P(b.W3s); w.W3Gs = P(b.W3Gs); w.W3Ts = P(b.W3Ts); w.W3GTs = P(b.W3GTs); w.W3JTs = P(b.W3JTs);
  w.b3 = params + g.boff[L];
  {
    int n1 = (dd + 1) * g.Hp[1];
    for (int l = 1; l <= L; ++l) n1 = g.Hp[l] > n1 ? g.Hp[l] : n1;
    tc_prep_first_kernel<<<(n1 + 255) / 256, 256, 0, st>>>(g, params, w);
    TC_KCK();
    if (L >= 2) {
      int nmax = 0;
      for (int l = 2; l <= L; ++l) nmax = g.Hp[l] * g.Hp[l - 1] > nmax ? g.Hp[l] * g.Hp[l - 1] : nmax;
      tc_prep_hidden_kernel<<<dim3((nmax + 255) / 256, L - 1), 256, 0, st>>>(g, params, w);
      TC_KCK();
    }
    tc_prep_last_kernel<<<(g.nout * HL + 255) / 256, 256, 0, st>>>(g, params, w);
    TC_KCK();
  }
  CUtensorMap mapW[TC_MAXL + 1], mapWT[TC_MAXL + 1], map3A, map3F, mapGA, mapT, mapGT, mapJT;
  for (int l = 2; l <= L; ++l) {
    if (!make_tmap_2d(&mapW[l], w.Ws[l], (uint64_t)2 * b.npad[l], g.Hp[l - 1], g.Hp[l - 1], HT)) return GPFLOW_ERR_CUDA;
    if (!make_tmap_2d(&mapWT[l], w.WTs[l], (uint64_t)2 * b.npadT[l], g.Hp[l], g.Hp[l], HT)) return GPFLOW_ERR_CUDA;
  }
  if (!make_tmap_2d(&map3A, w.W3s, (uint64_t)2 * g.rows3, HL, HL, MT)) return GPFLOW_ERR_CUDA;
  if (!make_tmap_2d(&map3F, w.W3s, (uint64_t)2 * g.rows3, HL, HL, HT)) return GPFLOW_ERR_CUDA;
  if (!make_tmap_2d(&mapGA, w.W3Gs, (uint64_t)2 * g.rowsG, HL, HL, MT)) return GPFLOW_ERR_CUDA;
  if (!make_tmap_2d(&mapT, w.W3Ts, (uint64_t)2 * g.HTp, g.ldT, g.ldT, HT)) return GPFLOW_ERR_CUDA;
  if (!make_tmap_2d(&mapGT, w.W3GTs, (uint64_t)nb * 2 * g.HTp, g.dpad, g.dpad, HT)) return GPFLOW_ERR_CUDA;
  if (!make_tmap_2d(&mapJT, w.W3JTs, (uint64_t)dd * 2 * g.HTp, g.nbpad, g.nbpad, HT)) return GPFLOW_ERR_CUDA;

  // GPFLOW_TC_GEN_T1=1: regenerate T_1[j] = sp_1 (.) W1x[:, j] inside the GEMM producers instead of materialising it.
  // Measured at C5 (2^14 samples): 990 ms per step against 948 ms with the materialised rows -- off by default.
  static const bool gen_t1 = [] { const char* e = getenv("GPFLOW_TC_GEN_T1"); return e != nullptr && e[0] == '1'; }();
  using LayerOp = OpLayer<HT, ST>;
  using LastOp = OpLast<2>;
  using LBOp = OpLB<HT, ST>;
  using W3Op = OpW3<HT, ST>;
  using WOp = OpW<HT, ST>;
  const int mt = S / MT;
  const int eb = 256;   // block of the element-wise kernels
  const int g_dS = (dd * S + eb - 1) / eb, g_nS = (nb * S + eb - 1) / eb, g_S = (S + eb - 1) / eb;

  // ---- one velocity evaluation, forward: activations of the slab at the stage point xpt, time t (up to SX / FX / SD) ----
  // 64-wide nets: tc_value_chain_kernel replaces tc_layer1_kernel and the value-row GEMM of every hidden layer
  static const bool chain_on = [] { const char* e = getenv("GPFLOW_TC_CHAIN"); return !(e && e[0] == '0'); }();   // A/B switch
  bool chain = chain_on && dd <= 16 && L >= 2 && L <= 4;
  for (int l = 1; l <= L; ++l) chain = chain && g.Hp[l] == 64;
  TcChainPtrs cq;
  memset(&cq, 0, sizeof(cq));
  for (int l = 1; l <= L; ++l) { cq.T[l] = P(b.T[l]); cq.sp[l] = P(b.sp[l]); cq.bp[l] = w.bp[l]; }
  if (chain) {
    static bool attr = false;
    if (!attr) {
      if (cudaFuncSetAttribute(tc_value_chain_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc_chain_smem(4)) != cudaSuccess) return GPFLOW_ERR_CUDA;
      attr = true;
    }
  }
  auto eval_forward = [&](const float* xpt, float t) -> int {
    if (chain)
      tc_value_chain_kernel<16><<<S / TC_CH_S, 256, tc_chain_smem(L), st>>>(g, params, w.W1T, cq, xpt, t, bnd, P(b.sj), (L == 1 || !gen_t1) ? 1 : 0);
    else if (dd <= 16)
      tc_layer1_kernel<16, 4><<<S / 4, g.Hp[1], 0, st>>>(g, w.W1T, w.bp[1], xpt, t, bnd, P(b.T[1]), P(b.sp[1]), P(b.sj), (L == 1 || !gen_t1) ? 1 : 0);
    else
      tc_layer1_kernel<TC_DMAX, TC_L1_SPB_WIDE><<<S / TC_L1_SPB_WIDE, g.Hp[1], 0, st>>>(g, w.W1T, w.bp[1], xpt, t, bnd, P(b.T[1]), P(b.sp[1]), P(b.sj), (L == 1 || !gen_t1) ? 1 : 0);
    TC_KCK();
    for (int l = 2; l <= L; ++l) {
      typename LayerOp::Params p;
      p.map = mapW[l];
      p.K = g.Hp[l - 1]; p.N = g.Hp[l]; p.lo_row = b.npad[l]; p.S = S; p.ntn = 1;
      p.bias = w.bp[l]; p.aux = P(b.sp[l]); p.out2 = P(b.sp[l]);
      p.gen_sp = nullptr; p.gen_w = nullptr;
      p.A = P(b.T[l - 1]) + (size_t)dd * S * p.K; p.M = S; p.mode = 0; p.out = P(b.T[l]) + (size_t)dd * S * p.N;
      if (!chain) TC_CK(launch_gemm<LayerOp>(p, mt, sm_count, st));
      p.A = P(b.T[l - 1]); p.M = (long long)dd * S; p.mode = 1; p.out = P(b.T[l]);
      if (l == 2 && gen_t1) { p.gen_sp = P(b.sp[1]); p.gen_w = w.W1T; }   // T_1 is not materialised
      TC_CK(launch_gemm<LayerOp>(p, dd * mt, sm_count, st));
    }
    tc_rsum_kernel<<<S, HL, 0, st>>>(g, P(b.T[L]), P(b.sj), P(b.R));
    TC_KCK();
    {   // F for every output row: one dense GEMM with column-layout stores
      typename LayerOp::Params p;
      p.map = map3F;
      p.K = HL; p.N = g.nout; p.lo_row = g.rows3; p.S = S; p.ntn = (g.nout + HT - 1) / HT;
      p.bias = w.b3; p.aux = nullptr; p.out2 = nullptr; p.gen_sp = nullptr; p.gen_w = nullptr;
      p.A = P(b.T[L]) + (size_t)dd * S * HL; p.M = S; p.mode = 3; p.out = P(b.F);
      TC_CK(launch_gemm<LayerOp>(p, mt * p.ntn, sm_count, st));
    }
    {
      typename LastOp::Params p;
      p.map3 = map3A; p.mapG = mapGA;
      p.TL = P(b.T[L]); p.R = P(b.R); p.G = P(b.G); p.D = P(b.D);
      p.d = dd; p.nb = nb; p.S = S; p.K = HL; p.nt = S / 256; p.lo3 = g.rows3; p.loG = g.rowsG;
      TC_CK(launch_gemm<LastOp>(p, (dd + nb) * p.nt, sm_count, st));
    }
    tc_vel1_kernel<<<g_nS, eb, 0, st>>>(g, xpt, P(b.sj), P(b.F), P(b.D), P(b.SX), P(b.FX), P(b.SD));
    TC_KCK();
    return GPFLOW_OK;
  };
  auto vel2 = [&](int sg, const float* xcur, float* ynext) -> int {
    tc_vel2_kernel<<<g_dS, eb, 0, st>>>(g, P(b.sj), P(b.F), P(b.G), P(b.SX), P(b.FX), P(b.SD), P(b.Q), P(b.v), sg, dt, xcur,
                                         P(b.ksum), ynext);
    TC_KCK();
    return GPFLOW_OK;
  };
  // ---- reverse sweep of the evaluation whose forward activations are in place; vbar / Qb set; result in ybar ----
  int wsplit_ns[TC_MAXL + 1] = {0};
  auto eval_backward = [&](const float* xpt, float t) -> int {
    tc_vb1_kernel<<<g_nS, eb, 0, st>>>(g, P(b.Qb), P(b.F), P(b.SQ), P(b.SXb));
    TC_KCK();
    tc_vb2_kernel<<<g_dS, eb, 0, st>>>(g, xpt, P(b.vbar), P(b.Q), P(b.F), P(b.D), P(b.SQ), P(b.SXb), P(b.ybar));
    TC_KCK();
    {
      typename LBOp::Params p;
      p.mapT = mapT; p.mapGT = mapGT; p.mapJT = mapJT;
      p.Qb = P(b.Qb); p.SX = P(b.SX); p.SQ = P(b.SQ); p.x = xpt; p.sj = P(b.sj);
      p.hbp = P(b.hbp); p.Rb = P(b.Rb); p.TbD = P(b.Ub[L]);
      p.d = dd; p.nb = nb; p.nout = g.nout; p.S = S; p.N = HL; p.HTp = g.HTp; p.dpad = g.dpad; p.mt = mt;
      p.nseg = b.nseg_hb; p.nchunk_p = b.nchunk_p;
      TC_CK(launch_gemm<LBOp>(p, (b.nseg_hb + nb + dd) * mt, sm_count, st));
    }
    if (b.w3_small) {
      W3SmallParams p;
      p.TL = P(b.T[L]); p.R = P(b.R); p.Qb = P(b.Qb); p.SX = P(b.SX); p.SQ = P(b.SQ); p.x = xpt; p.sj = P(b.sj);
      p.gW3 = P(b.gW3);
      p.d = dd; p.nb = nb; p.nout = g.nout; p.S = S;
      p.nkept = 0;
      for (int i = 0; i < dd; ++i) p.nkept += (i < nb - 1 ? i : nb - 1) + 1;
      p.nF = (p.nkept + W3S_ROWS - 1) / W3S_ROWS; p.nDt = (nb + W3S_ROWS - 1) / W3S_ROWS; p.nGt = (dd + W3S_ROWS - 1) / W3S_ROWS;
      p.nsplit = b.w3_nsplit;
      p.spb = ((S + p.nsplit - 1) / p.nsplit + W3S_CHUNK - 1) / W3S_CHUNK * W3S_CHUNK;
      tc_w3_small_kernel<<<dim3(p.nF + dd * p.nDt + nb * p.nGt, p.nsplit), 256, 0, st>>>(p);
      TC_KCK();
    } else {
      typename W3Op::Params p;
      p.TL = P(b.T[L]); p.R = P(b.R); p.Qb = P(b.Qb); p.SX = P(b.SX); p.SQ = P(b.SQ); p.x = xpt; p.sj = P(b.sj);
      p.gW3 = P(b.gW3);
      p.d = dd; p.nb = nb; p.nout = g.nout; p.S = S; p.K = HL;
      p.nF = (g.nout + MT - 1) / MT;
      p.n_out = p.nF + dd + nb;
      p.nseg = S / (KC * TC_SEG);
      p.nsplit = b.w3_nsplit;
      p.items = p.n_out * p.nsplit;
      p.G = p.items < sm_count ? p.items : sm_count;
      p.segs_per_item = (p.nseg + p.nsplit - 1) / p.nsplit;
      p.rounds = (p.items + p.G - 1) / p.G * p.segs_per_item;
      // the grid must be exactly G so that every work item stays with one CTA
      using SMW = GemmSmem<W3Op>;
      static bool attr = false;
      if (!attr) { TC_CK(cudaFuncSetAttribute(tc_gemm_kernel<W3Op>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMW::kBytes)); attr = true; }
      tc_gemm_kernel<W3Op><<<p.G, kThreads, SMW::kBytes, st>>>(p);
      TC_KCK();
    }
    tc_db3_kernel<<<dim3((nb + 15) / 16, (dd + 15) / 16, S / 64), dim3(16, 16), 0, st>>>(g, P(b.Qb), xpt, P(b.SX), P(b.SQ), 64, P(b.zb3));
    TC_KCK();
    tc_tb_last_kernel<<<S, HL, 0, st>>>(g, P(b.T[L]), P(b.sp[L]), P(b.Rb), P(b.hbp), b.nseg_hb, P(b.sj), xpt, P(b.Ub[L]),
                                        P(b.ybar), L == 1 ? 1 : 0);
    TC_KCK();
    for (int l = L; l >= 2; --l) {
      {
        typename WOp::Params p;
        p.Ub = P(b.Ub[l]); p.Tm = P(b.T[l - 1]); p.gWp = P(b.gWp[l]);
        p.gen_sp = (l == 2 && gen_t1) ? P(b.sp[1]) : nullptr; p.gen_w = w.W1T; p.S = S; p.d = dd;
        p.Ml = g.Hp[l]; p.Nl = g.Hp[l - 1];
        const int mtl = (p.Ml + MT - 1) / MT;
        p.nchunks = (dd + 1) * S / KC;
        p.nsplit = sm_count / mtl < 1 ? 1 : sm_count / mtl;
        if (p.nsplit > p.nchunks) p.nsplit = p.nchunks;
        p.G = mtl * p.nsplit;
        p.cps = (p.nchunks + p.nsplit - 1) / p.nsplit;
        p.nseg = (p.cps + TC_SEG - 1) / TC_SEG;
        wsplit_ns[l] = p.nsplit;
        using SMW = GemmSmem<WOp>;
        static bool attr = false;
        if (!attr) { TC_CK(cudaFuncSetAttribute(tc_gemm_kernel<WOp>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMW::kBytes)); attr = true; }
        tc_gemm_kernel<WOp><<<p.G, kThreads, SMW::kBytes, st>>>(p);
        TC_KCK();
      }
      tc_colsum_kernel<<<dim3((g.Hp[l] + 31) / 32, b.nz), dim3(32, 8), 0, st>>>(S, g.Hp[l], b.s_per_z,
                                                                                P(b.Ub[l]) + (size_t)dd * S * g.Hp[l], P(b.zb[l]));
      TC_KCK();
      {
        typename LayerOp::Params p;
        p.map = mapWT[l];
        p.K = g.Hp[l]; p.N = g.Hp[l - 1]; p.lo_row = b.npadT[l]; p.S = S; p.ntn = 1;
        p.bias = nullptr; p.aux = nullptr; p.out2 = nullptr; p.gen_sp = nullptr; p.gen_w = nullptr;
        p.A = P(b.Ub[l]); p.M = (long long)(dd + 1) * S; p.mode = 2; p.out = P(b.Ub[l - 1]);
        TC_CK(launch_gemm<LayerOp>(p, (dd + 1) * mt, sm_count, st));
      }
      if (l - 1 >= 2) {
        tc_tb_mid_kernel<<<S, g.Hp[l - 1], 0, st>>>(g, l - 1, P(b.T[l - 1]), P(b.sp[l - 1]), P(b.Ub[l - 1]));
        TC_KCK();
      }
    }
    tc_layer1_bwd_a_kernel<<<S / TC_L1B_SPB, g.Hp[1], 0, st>>>(g, w.W1T, P(b.T[1]), P(b.sp[1]), P(b.Ub[1]), P(b.ybar));
    TC_KCK();
    if (b.l1b_small)
      tc_layer1_bwd_b_small_kernel<16><<<dim3(S / TC_L1B_CH, (dd + 15) / 16, g.Hp[1] / 64), 256, 0, st>>>(g, P(b.sp[1]), P(b.Ub[1]), xpt, t, P(b.zW1));
    else
      tc_layer1_bwd_b_kernel<<<dim3(g.Hp[1] / 32, (dd + 7) / 8, b.nz), dim3(32, 8), 0, st>>>(g, P(b.sp[1]), P(b.Ub[1]), xpt, t,
                                                                                         b.s_per_z, P(b.zW1));
    TC_KCK();
    return GPFLOW_OK;
  };

  const size_t dS = (size_t)dd * S;
  auto slot = [&](int i) { return P(b.stash) + (size_t)i * dS; };
  int rc;
  for (long long b0 = 0; b0 < B; b0 += S) {
    const int nvalid = (int)(B - b0 < S ? B - b0 : S);
    tc_prologue_kernel<<<g_dS, eb, 0, st>>>(g, x_in, b0, nvalid, (wrap && v_out == nullptr) ? 1 : 0, sq2s, slot(0));
    TC_KCK();
    if (v_out != nullptr) {   // VelocityField.__call__(x, t)
      if ((rc = eval_forward(slot(0), (float)t_eval)) != GPFLOW_OK) return rc;
      if ((rc = vel2(-1, nullptr, nullptr)) != GPFLOW_OK) return rc;
      tc_export_kernel<<<g_dS, eb, 0, st>>>(g, P(b.v), b0, nvalid, v_out);
      TC_KCK();
      continue;
    }
    // ---------------- forward time loop ----------------
    float t = t0;
    for (int i = 0; i < N; ++i) {
      if (traj_out != nullptr) {
        tc_export_kernel<<<g_dS, eb, 0, st>>>(g, slot(4 * i), b0, nvalid, traj_out + (size_t)i * B * dd);
        TC_KCK();
      }
      for (int sg = 0; sg < 4; ++sg) {
        const float ts = sg == 0 ? t : (sg == 3 ? t + dt : t + hdt);
        if ((rc = eval_forward(slot(4 * i + sg), ts)) != GPFLOW_OK) return rc;
        if ((rc = vel2(sg, slot(4 * i), slot(4 * i + sg + 1))) != GPFLOW_OK) return rc;
      }
      if (want_lagr) {   // 1/2 |v(x_{i+1}, t_i)|^2  (res_net.py:259-262,292-294)
        if ((rc = eval_forward(slot(4 * i + 4), t)) != GPFLOW_OK) return rc;
        if ((rc = vel2(-1, nullptr, nullptr)) != GPFLOW_OK) return rc;
        tc_lagr_kernel<<<g_S, eb, 0, st>>>(g, P(b.v), nvalid, i == 0 ? 1 : 0, P(b.lg_s));
        TC_KCK();
      }
      t = t + dt;
    }
    if (traj_out != nullptr) {
      tc_export_kernel<<<g_dS, eb, 0, st>>>(g, slot(4 * N), b0, nvalid, traj_out + (size_t)N * B * dd);
      TC_KCK();
    }
    if (want_lagr) { tc_sum_kernel<<<1, 256, 0, st>>>(S, P(b.lg_s), P(b.scal) + 2); TC_KCK(); }
    const int mode = op == GPFLOW_OP_FORWARD ? 0 : (op == GPFLOW_OP_TRAIN_STEP ? 1 : 2);
    tc_epilogue_kernel<<<g_S, eb, 0, st>>>(g, slot(4 * N), x_in, x_data, grad_y, b0, nvalid, wrap ? 1 : 0, sigma, (float)d->loss_scale,
                                           mode, y_out, wrap ? logdet_out : nullptr, train ? P(b.xb) : nullptr, P(b.ot_s));
    TC_KCK();
    if (!train) continue;
    if (mode == 1) { tc_sum_kernel<<<1, 256, 0, st>>>(S, P(b.ot_s), P(b.scal)); TC_KCK(); }
    // ---------------- reverse time loop (A.5): the stage points were kept, activations are recomputed ----------------
    for (int i = N - 1; i >= 0; --i) {
      float ti = t0;
      for (int q = 0; q < i; ++q) ti = ti + dt;
      for (int sg = 3; sg >= 0; --sg) {
        const float ts = sg == 0 ? ti : (sg == 3 ? ti + dt : ti + hdt);
        const float wk = (sg == 1 || sg == 2) ? 2.f * c6 : c6;
        const float ak = sg == 3 ? 0.f : (sg == 2 ? dt : hdt);
        const float* xpt = slot(4 * i + sg);
        if ((rc = eval_forward(xpt, ts)) != GPFLOW_OK) return rc;
        if ((rc = vel2(-1, nullptr, nullptr)) != GPFLOW_OK) return rc;   // Q of the stage (needed by the reverse of A.2)
        tc_vbar_kernel<<<g_dS, eb, 0, st>>>(g, wk, ak, P(b.xb), P(b.ybprev), P(b.sj), P(b.vbar), P(b.Qb));
        TC_KCK();
        if ((rc = eval_backward(xpt, ts)) != GPFLOW_OK) return rc;
        tc_stage_bwd_kernel<<<g_dS, eb, 0, st>>>(g, sg, P(b.ybar), P(b.ybsum), P(b.ybprev), P(b.xb));
        TC_KCK();
      }
    }
    if (grad_x_out != nullptr) {
      tc_gradx_kernel<<<g_dS, eb, 0, st>>>(g, P(b.xb), x_in, b0, nvalid, wrap ? 1 : 0, sigma, grad_x_out);
      TC_KCK();
    }
  }
  if (v_out != nullptr) return GPFLOW_OK;
  // ---- finalize: padded accumulators -> the caller's flat gradient, scalars ----
  if (scalars_out != nullptr || grads_out != nullptr) {
    TcTabs tabs;
    memset(&tabs, 0, sizeof(tabs));
    if (train)
    {
      // the per-split planes of the bias / first-layer gradients, summed once per call in a fixed order
      const int n1 = g.Hp[1] * (dd + 2);
      tc_reduce_add_kernel<<<(n1 + 255) / 256, 256, 0, st>>>(b.nz1, n1, P(b.zW1), P(b.gW1));
      TC_KCK();
      tc_reduce_add_kernel<<<(g.nout + 255) / 256, 256, 0, st>>>(S / 64, g.nout, P(b.zb3), P(b.gb3));
      TC_KCK();
      for (int l = 2; l <= L; ++l) {
        tc_reduce_add_kernel<<<(g.Hp[l] + 255) / 256, 256, 0, st>>>(b.nz, g.Hp[l], P(b.zb[l]), P(b.gb[l]));
        TC_KCK();
      }
      for (int l = 2; l <= L; ++l) { tabs.gWp[l] = P(b.gWp[l]); tabs.gb[l] = P(b.gb[l]); tabs.ns[l] = wsplit_ns[l]; }
      tabs.ns3 = 3 * b.w3_nsplit;
    }
    const int count = (train && grads_out != nullptr) ? g.P : 3;
    tc_finalize_kernel<<<(count + 255) / 256, 256, 0, st>>>(
        g, train ? P(b.gW1) : nullptr, tabs, train ? P(b.gW3) : nullptr, train ? P(b.gb3) : nullptr, P(b.scal),
        (float)d->loss_scale, (float)(d->dt_constraint * d->loss_scale), train ? grads_out : nullptr, scalars_out);
    TC_KCK();
  }
  return GPFLOW_OK;
}

}  // namespace tc
}  // namespace gpflow
