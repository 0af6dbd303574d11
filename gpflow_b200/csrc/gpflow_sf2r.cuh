// gpflow_b200 -- "sf2r": second formulation of the 2-D stream-function train-step kernel (sm_100a, fp32, d = 2,
// nb_div_blocks = 1, two hidden layers of width <= 15).  Same reference call sites, same results and the same thread
// mapping as gpflow_sf2.cuh (one thread = one sample, one warp = a tile of 32 samples, TMA-streamed stash); what changes
// is HOW the gradient of the stream function is obtained.
//
// sf2 follows the reference literally: forward-mode tangents through the network (vector_field.py:92-128), i.e. per
// evaluation three 15x15 contractions forward (a2, U2[:,0], U2[:,1]), three back through W2 and a rank-3 update of dW2.
// For d = 2, nb = 1 the velocity only needs the scalar psi = f_0 - f_1 and its gradient (gpflow_sf2.cuh header), and the
// gradient of ONE scalar is cheaper in reverse mode:
//     a1 = W1 [x; t] + b1, h1 = tanh a1, sp1 = 1 - h1^2;   a2 = W2 h1 + b2, h2 = tanh a2, sp2 = 1 - h2^2
//     psi = w3 . h2 + b3;   ws = w3 (.) sp2;   z = W2^T ws;   q = sp1 (.) z;   d_j psi = sum_m W1[m][j] q[m]
// -- two contractions instead of three.  Its own reverse sweep (cotangents psib, gb_j of psi, d_j psi) is
//     c  = sp1 (.) (W1x gb)            tb = W2 c                 e  = psib h2 + sp2 (.) tb           -> dw3
//     ab2 = ws (.) (psib - 2 h2 (.) tb)                          hb = W2^T ab2,   z = W2^T ws (recomputed, same rows)
//     dW2 += ab2 h1^T + ws c^T   (rank 2 per sample)             db2 += ab2
//     ab1 = sp1 (.) hb - 2 h1 (.) z (.) c                        dW1 += ab1 [x; t]^T + q gb^T,   db1 += ab1
//     xbar = W1x^T ab1 + (closed-form part)
// -- three contractions that share ONE pass over the rows of W2, and a rank-2 update: 7 contraction-equivalents per
// evaluation instead of 9, and the stash shrinks from 72 to 32 floats per sample and evaluation (x, h1, h2: 128 B, i.e.
// 15.4 KB per sample and train step instead of 33.8 KB).  The arithmetic is plain fp32 FMA as before; only the order of
// the sums differs from sf2 / the reference (both agree with the fp64 oracle to ~1e-6).
//
// Shared-memory layout per warp:
//   stash buffers  2 x [32 samples][32 floats], 128-byte rows whose 16-byte chunks are XOR-swizzled with (sample & 7):
//                  chunks 0-3 = (x0, h1[0..14]), chunks 4-7 = (x1, h2[0..14]).  A row is exactly what goes to HBM (one
//                  4 KB bulk copy per evaluation); the swizzle makes the per-lane 128-bit accesses conflict free
//                  without padding bytes in the stash.
//   work rows      [32 samples][68 floats] (reverse sweep only): ab2 | ws | c | e | (x0 x1 gb0 gb1), later ab1 | q over
//                  ab2 | ws.  Left/right factors of the two rank updates, read by the other lanes of the warp.
// "Position" p = unit + 1 indexes the 16-wide rows (position 0 is the x slot of the stash row): it doubles as the bias
// column of layer 2 in the forward contraction and in the dW2 tile.
#pragma once
#include "gpflow_sf2.cuh"

namespace gpflow {

template <int H1_, int H2_, int NW_>
struct SF2R {
  static constexpr int H1 = H1_, H2 = H2_, NW = NW_;
  static_assert(H1 >= 1 && H1 <= 15 && H2 >= 1 && H2 <= 15, "position 0 of the 16-wide rows is the x / bias slot");
  typedef SF2<H1_, H2_, NW_> Flat;  // flat parameter offsets
  static constexpr int NPARAM = Flat::NPARAM;
  // packed weights (floats)
  static constexpr int W1P = 0;          // [16][4]  unit m: {W1[m][0], W1[m][1], W1[m][2] (time), b1[m]}; rows >= H1 zero
  static constexpr int W2F = W1P + 64;   // [16][16] forward rows AND columns by position: W2F[0][k+1] = b2[k],
                                         //          W2F[m+1][k+1] = W2[k][m], column 0 = 0
  static constexpr int W2B = W2F + 256;  // [16][16] backward rows AND columns by position: W2B[k+1][m+1] = W2[k][m], row 0 =
                                         //          column 0 = 0 (register pairs line up with the position-indexed rows)
  static constexpr int W3P = W2B + 256;  // [16]     by position: W3P[0] = 0, W3P[k+1] = W3[0][k] - W3[1][k]
  static constexpr int B3C = W3P + 16;   // [4]      b3[0] - b3[1], 0, 0, 0
  static constexpr int W1Q = B3C + 4;    // [8][8]   layer 1 by POSITION PAIRS (2j, 2j+1): {W1[.][0] x2, W1[.][1] x2, W1[.][2] x2, b1 x2};
                                         //          position 0 and positions beyond H1 are zero (operands of packed fp32x2 FMAs)
  static constexpr int WEND = W1Q + 64;
  static constexpr int WPACK = 704;      // floats of the packed-weight staging buffer / constant array (>= WEND)
  static_assert(WEND <= WPACK, "packed weights do not fit the constant array");
  // per-warp buffers (floats)
  static constexpr int SROW = 32, S_FL = 32 * SROW, STASH_BYTES = S_FL * 4;   // one evaluation of one tile: 4 KB
  static constexpr int WS = 68;                                               // work row, 17 x 16 B (odd: conflict free)
  static constexpr int O_AB = 0, O_WS = 16, O_C = 32, O_E = 48, O_XG = 64;
  static constexpr int WK_FL = 32 * WS;
  static constexpr int WARP_FL = 2 * S_FL + WK_FL;
  // second-level accumulators: 16 dW2 tile | 4 dw3 | db3 | 16 dW1 tile | ot
  static constexpr int G_W2 = 0, G_W3 = 16, G_B3 = 20, G_W1 = 21, G_OT = 37, NACC = 38;
  static constexpr int SCT = 16;

  static size_t smem_bytes(int n_steps) {
    return (size_t)(WEND + SCT + round_up(n_steps + 1, 4) + round_up(NPARAM, 4) + NW * WARP_FL) * sizeof(float) + 128;
  }
};

// packed weight i from the flat nn.Linear parameters (vector_field.py:56-61)
template <class K>
GPF_DEV float sf2r_pack_value(const float* __restrict__ raw, int i) {
  typedef typename K::Flat F;
  constexpr int H1 = K::H1, H2 = K::H2;
  float v = 0.0f;
  if (i < K::W2F) {
    const int m = i >> 2, c = i & 3;
    if (m < H1) v = c < 3 ? raw[F::OFF_W1 + 3 * m + c] : raw[F::OFF_B1 + m];
  } else if (i < K::W2B) {
    const int p = (i - K::W2F) >> 4, pk = (i - K::W2F) & 15;
    if (pk >= 1 && pk - 1 < H2)
      v = p == 0 ? raw[F::OFF_B2 + pk - 1] : (p - 1 < H1 ? raw[F::OFF_W2 + (pk - 1) * H1 + (p - 1)] : 0.0f);
  } else if (i < K::W3P) {
    const int p = (i - K::W2B) >> 4, pm = (i - K::W2B) & 15;
    if (p >= 1 && p - 1 < H2 && pm >= 1 && pm - 1 < H1) v = raw[F::OFF_W2 + (p - 1) * H1 + (pm - 1)];
  } else if (i < K::B3C) {
    const int p = i - K::W3P;
    if (p >= 1 && p - 1 < H2) v = raw[F::OFF_W3 + (p - 1)] - raw[F::OFF_W3 + H2 + (p - 1)];
  } else if (i == K::B3C) {
    v = raw[F::OFF_B3] - raw[F::OFF_B3 + 1];
  } else if (i >= K::W1Q && i < K::W1Q + 64) {
    const int j = (i - K::W1Q) >> 3, e = (i - K::W1Q) & 7, pos = 2 * j + (e & 1), c = e >> 1;   // c: x0, x1, t, bias
    if (pos >= 1 && pos - 1 < H1) v = c < 3 ? raw[F::OFF_W1 + 3 * (pos - 1) + c] : raw[F::OFF_B1 + pos - 1];
  }
  return v;
}

// Where the (warp-uniform) packed weights are read from.  A warp-uniform LDS.128 costs 2 shared-memory wavefronts for 16
// useful bytes (measured, profiles/r2_sf2r_*), and with 7 instead of 9 contractions per evaluation the shared-memory pipe
// became the limiter (84 % busy); the constant bank serves the same loads through the uniform datapath (LDCU / LDC).
// GPF_SF2R_CONST is a bit mask: 1 = forward evaluation, 2 = the rolled W2 pass of the reverse sweep, 4 = the other
// weight loads of the reverse sweep.  The constant copy is per device and shared by all calls: see launch_sf2r.
#ifndef GPF_SF2R_CONST
#define GPF_SF2R_CONST 1   // measured at C4 (tools/sf2r_const_ab.sh): 0: 1.34e8, 1: 1.48e8, 3: 1.47e8, 7: 1.40e8 samples/s
#endif
__constant__ float sf2r_cw[704];
// GPF_SF2R_MMA: 0 (default) = the rank-2 dW2 update of the reverse sweep as FFMA2 register tiles; 1 / 2 = on the tensor
// cores (mma.sync m16n8k8, 3xTF32; split by cvt.rna / by truncation).  Parity-green and measured SLOWER at C4
// (tools/sf2r_mma_ab.sh, same box): 111.6 ms (0), 121.8 ms (1), 114.4 ms (2) -- the FMA pipe drops from 67 % to 59 % busy but
// the 64 fragment loads, 128 split operations and 48 mma per evaluation issue as many instructions as the 64 LDS.128 +
// 256 FFMA2 they replace, and the kernel is issue-bound as much as FMA-bound (issue slots 67 %).
#ifndef GPF_SF2R_MMA
#define GPF_SF2R_MMA 0
#endif
// GPF_SF2R_TANH: 0 = tanh_real (<= 2 ulp relative, default), 1 = tanh_exp2 (absolute accuracy ~3e-7).  Measured at C4
// (tools/sf2r_tanh_ab.sh): 1.46e8 vs 1.60e8 samples/s and the same errors against the fp64 oracle at weights N(0, 0.1..0.6^2),
// but at the reference's own initialisation (weights N(0, 1e-3^2): |a| ~ 1e-3, where tanh x ~ x needs RELATIVE accuracy)
// the 12-step training replays of the unmodified reference fail their 1e-5 bar -- so the faster variant is not used.
#ifndef GPF_SF2R_TANH
#define GPF_SF2R_TANH 0
#endif
GPF_DEV float sf2r_tanh(float x) { return GPF_SF2R_TANH ? tanh_exp2(x) : tanh_real(x); }
GPF_DEV float2 sf2r_tanh2(float2 x) { return GPF_SF2R_TANH ? make_float2(tanh_exp2(x.x), tanh_exp2(x.y)) : tanh_real2(x); }
template <int BIT>
GPF_DEV float4 ldwr4(const float* W, int off) {
  if (GPF_SF2R_CONST & BIT) return *reinterpret_cast<const float4*>(&sf2r_cw[off]);
  return ldsw4(W + off);
}

// chunk c (0..7) of a swizzled stash row; l7 = sample & 7
GPF_DEV int sw_chunk(int c, int l7) { return (c ^ l7) << 2; }

template <int BIT>
GPF_DEV void load_w16(const float* W, int off, float2 (&wp)[8]) {
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float4 t = ldwr4<BIT>(W, off + 4 * q);
    wp[2 * q] = make_float2(t.x, t.y);
    wp[2 * q + 1] = make_float2(t.z, t.w);
  }
}

// ---------------------------------------------------------------------------------------------------------
// forward: v(x, ts).  Writes the stash row (x0, h1 | x1, h2) of this lane's sample.
// ---------------------------------------------------------------------------------------------------------
template <class K, bool STASH = true>
GPF_DEV void sf2r_forward_eval(const float* __restrict__ W, float* __restrict__ Srow, int l7, float x0, float x1, float ts,
                               float bnd, float& v0, float& v1) {
  constexpr int H1 = K::H1, H2 = K::H2;
  // Everything is held as POSITION PAIRS (2j, 2j+1), position p = unit + 1, so that the element-wise stages are packed
  // fp32x2 instructions too and the pairs coincide with the 128-bit stash chunks (no register shuffles).
  // ---- layer 1: h1 = tanh(W1 [x; t] + b1)   (vector_field.py:75-90)
  float2 hpp[8];   // hpp[0].x = x0 (position 0), then h1
  const float2 x0d = make_float2(x0, x0), x1d = make_float2(x1, x1), tsd = make_float2(ts, ts), one2 = make_float2(1.0f, 1.0f);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float4 w0 = ldwr4<1>(W, K::W1Q + 8 * j), w1 = ldwr4<1>(W, K::W1Q + 8 * j + 4);
    const float2 a = __ffma2_rn(make_float2(w0.x, w0.y), x0d, __ffma2_rn(make_float2(w0.z, w0.w), x1d,
                                __ffma2_rn(make_float2(w1.x, w1.y), tsd, make_float2(w1.z, w1.w))));
    hpp[j] = sf2r_tanh2(a);   // zero weights -> tanh(0) = 0 at the unused positions
  }
  hpp[0].x = x0;
  if (STASH) {
#pragma unroll
    for (int c = 0; c < 4; ++c) sts4(Srow + sw_chunk(c, l7), hpp[2 * c].x, hpp[2 * c].y, hpp[2 * c + 1].x, hpp[2 * c + 1].y);
  }
  // ---- layer 2 (position p unrolled): a2 = b2 + W2 h1, two output positions per packed FMA
  float2 a2p[8];
  load_w16<1>(W, K::W2F, a2p);  // position 0: the bias row times 1
#pragma unroll
  for (int p = 1; p <= H1; ++p) {
    float2 wp[8];
    load_w16<1>(W, K::W2F + 16 * p, wp);
    const float h = (p & 1) ? hpp[p >> 1].y : hpp[p >> 1].x;
    const float2 hd = make_float2(h, h);
#pragma unroll
    for (int j = 0; j < 8; ++j) a2p[j] = __ffma2_rn(wp[j], hd, a2p[j]);
  }
  // ---- activation 2, collapsed output row: psi = w3 . h2 + b3, ws = w3 (.) sp2   (w3 = 0 at position 0)
  float2 h2pp[8], ws2[8], psip = make_float2(ldwr4<1>(W, K::B3C).x, 0.0f);
#pragma unroll
  for (int j = 0; j < 8; ++j) h2pp[j] = sf2r_tanh2(a2p[j]);
  h2pp[0].x = x1;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const float4 w3v = ldwr4<1>(W, K::W3P + 4 * c);
    const float2 w3a = make_float2(w3v.x, w3v.y), w3b = make_float2(w3v.z, w3v.w);
    const float2 spa = __ffma2_rn(make_float2(-h2pp[2 * c].x, -h2pp[2 * c].y), h2pp[2 * c], one2);
    const float2 spb = __ffma2_rn(make_float2(-h2pp[2 * c + 1].x, -h2pp[2 * c + 1].y), h2pp[2 * c + 1], one2);
    psip = __ffma2_rn(w3a, h2pp[2 * c], psip);
    psip = __ffma2_rn(w3b, h2pp[2 * c + 1], psip);
    ws2[2 * c] = __fmul2_rn(w3a, spa);
    ws2[2 * c + 1] = __fmul2_rn(w3b, spb);
    if (STASH) sts4(Srow + sw_chunk(4 + c, l7), h2pp[2 * c].x, h2pp[2 * c].y, h2pp[2 * c + 1].x, h2pp[2 * c + 1].y);
  }
  const float psi = psip.x + psip.y;
  // ---- z = W2^T ws (position p of h2 unrolled); zp by position of h1
  float2 zp[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) zp[j] = make_float2(0.0f, 0.0f);
#pragma unroll
  for (int p = 1; p <= H2; ++p) {
    float2 wp[8];
    load_w16<1>(W, K::W2B + 16 * p, wp);
    const float w = (p & 1) ? ws2[p >> 1].y : ws2[p >> 1].x;
    const float2 wd = make_float2(w, w);
#pragma unroll
    for (int j = 0; j < 8; ++j) zp[j] = __ffma2_rn(wp[j], wd, zp[j]);
  }
  // ---- grad psi = W1x^T (sp1 (.) z)   (position 0: zero weights and z = 0)
  float2 g0p = make_float2(0.0f, 0.0f), g1p = g0p;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float4 w0 = ldwr4<1>(W, K::W1Q + 8 * j);
    const float2 q = __fmul2_rn(__ffma2_rn(make_float2(-hpp[j].x, -hpp[j].y), hpp[j], one2), zp[j]);
    g0p = __ffma2_rn(make_float2(w0.x, w0.y), q, g0p);
    g1p = __ffma2_rn(make_float2(w0.z, w0.w), q, g1p);
  }
  const float g0 = g0p.x + g0p.y, g1 = g1p.x + g1p.y;
  // ---- closed form A.2 for d = 2, nb = 1 (velocity_field.py:20-79)
  const float s0 = fmaf(x0, x0, -bnd), s1 = fmaf(x1, x1, -bnd);
  const float q0 = fmaf(2.0f * psi, x1, s1 * g1);
  const float q1 = -fmaf(2.0f * psi, x0, s0 * g0);
  v0 = s0 * q0;
  v1 = s1 * q1;
}

}  // namespace gpflow
