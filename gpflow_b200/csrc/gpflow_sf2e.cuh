// gpflow_b200 -- "sf2e": the Euler penalization (res_net.py:319-358, velocity_field.py:180-185) on the sf2r structure
// (gpflow_sf2r.cuh): d = 2, nb_div_blocks = 1, two hidden layers of width <= 15, fp32.
//
//   pen = mean_b (z . J_a w - w . J_a z)^2,   a(x) = [v(x + delta v(x, t), t + delta) - v(x, t)] / delta,   delta = 2 sqrt(eps)
//
// The two functorch.vjp calls of the reference are restated as two JVPs (oracle A.7): J_a w is the epsilon part of
// a(x + eps w), obtained by evaluating the velocity in DUAL arithmetic (value, epsilon part), and the training gradient of
// that epsilon part is the epsilon part of the same reverse sweep run in dual arithmetic (reverse-over-forward; this is
// where tanh'' and tanh''' appear without being derived by hand) -- exactly what gpflow_small.cuh does with Dual<T>.
// Here the dual number IS the fp32x2 register pair: a weight is a plain scalar, so one packed FMA (FFMA2 with the
// scalar operand broadcast) advances the value and the epsilon part of one unit at once.  The formulas are those of
// sf2r_forward_eval / sf2r_reverse_eval with every product replaced by the dual product.
//
// Per RK4 step the penalty costs 4 dual evaluations forward (2 directions x 2 points) and 4 dual reverse sweeps; their
// rows -- x, h1, h2 as (value, epsilon) pairs, 64 floats = 256 B per sample -- go through the same TMA-streamed stash.
#pragma once
#include "gpflow_sf2r_rev.cuh"

namespace gpflow {

// contraction loops of the dual evaluations: unroll factor, and where their weight rows come from (a bit of GPF_SF2R_CONST:
// 1 = constant bank like the unrolled plain forward evaluation, 2 = shared memory).  Measured on the C4 Euler row.  With
// free-running warps (first build) unroll 1 / 3 / 5 and either source were all 157.6-159.2 ms -- every warp missed the
// instruction cache on its own whatever the code size.  Since the CTA's warps are kept in phase (gpflow_sf2e_kernel.cuh,
// cta_align) they share the instruction stream and the unrolled form pays (tools/sf2e_ab2.sh, sf2e_ab3.sh,
// profiles/r2_sf2e_aligned_ab.txt): unroll 1 / 3 / 5 / 15 from shared memory 103.2 / 101.6 / 101.2 / 97.9 ms, unroll 15 from
// the constant bank 95.9 ms (4.37e7 samples/s).
#ifndef GPF_SF2E_UNROLL
#define GPF_SF2E_UNROLL 15
#endif
#ifndef GPF_SF2E_WBIT
#define GPF_SF2E_WBIT 1
#endif
#ifndef GPF_SF2E_W1BIT
#define GPF_SF2E_W1BIT 4   // layer-1 weight rows and b3 in the dual reverse sweep: 4 = shared memory (89.8 ms), 1 = constant bank (99.7 ms; tools/sf2e_ab6.sh)
#endif
#ifndef GPF_SF2E_RBIT
#define GPF_SF2E_RBIT 1   // weight rows of the dual REVERSE contraction: 2 = shared memory (91.65 ms), 1 = constant bank (89.70 ms; tools/sf2e_ab5.sh)
#endif
constexpr int kSF2EUnroll = GPF_SF2E_UNROLL;
// pair-accumulator rank update of sf2e_reverse_eval (14 % of the instructions, top stall short_scoreboard on its LDS.128): unroll
// 2 / 4 / 16 -> 96.1 / 96.4 / 91.7 ms on the C4 Euler row (tools/sf2e_ab4.sh; 56 B of spills at 4 and 16)
#ifndef GPF_SF2E_RANK_UNROLL
#define GPF_SF2E_RANK_UNROLL 16
#endif
constexpr int kSF2ERankUnroll = GPF_SF2E_RANK_UNROLL;   // pair-accumulator rank update of sf2e_reverse_eval
typedef float2 du;  // dual number: .x = value, .y = epsilon part
// GPF_SF2E_PACKED = 1: the element-wise dual operations as packed fp32x2 instructions as well (the scalar forms are 17 % FMUL +
// FADD of the executed instructions, profiles/r2_sf2e_ncu_summary.txt).  Measured on the C4 Euler row, same box, alternating
// builds (tools/sf2e_packed_ab.sh, profiles/r2_sf2e_packed_ab.txt): 152.8 / 155.9 ms packed against 139.4 / 140.7 ms scalar --
// a packed instruction holds the FMA pipe for two cycles and needs aligned register pairs (extra moves at 255 registers),
// so fewer instructions do not pay here (with aligned warps and unrolled loops: 98.3 against 97.9 ms, still no gain).  Default 0.
#ifndef GPF_SF2E_PACKED
#define GPF_SF2E_PACKED 0
#endif
GPF_DEV du d_make(float v, float e) { return make_float2(v, e); }
GPF_DEV du d_fma(float s, du a, du c) { return __ffma2_rn(make_float2(s, s), a, c); }   // s a + c, s a plain scalar
#if GPF_SF2E_PACKED
GPF_DEV du d_add(du a, du b) { return __fadd2_rn(a, b); }
GPF_DEV du d_sub(du a, du b) { return __ffma2_rn(make_float2(-1.0f, -1.0f), b, a); }
GPF_DEV du d_neg(du a) { return make_float2(-a.x, -a.y); }
GPF_DEV du d_mul(du a, du b) { return __ffma2_rn(make_float2(a.x, a.x), b, make_float2(0.0f, a.y * b.x)); }
GPF_DEV du d_scale(float s, du a) { return __fmul2_rn(make_float2(s, s), a); }
GPF_DEV du d_sp(du h) { return __ffma2_rn(make_float2(-h.x, -h.x), make_float2(h.x, 2.0f * h.y), make_float2(1.0f, 0.0f)); }  // 1 - h^2
GPF_DEV du d_sq_minus(du x, float b) { return __ffma2_rn(make_float2(x.x, x.x), make_float2(x.x, 2.0f * x.y), make_float2(-b, 0.0f)); }  // x^2 - b
#else
GPF_DEV du d_add(du a, du b) { return make_float2(a.x + b.x, a.y + b.y); }
GPF_DEV du d_sub(du a, du b) { return make_float2(a.x - b.x, a.y - b.y); }
GPF_DEV du d_neg(du a) { return make_float2(-a.x, -a.y); }
GPF_DEV du d_mul(du a, du b) { return make_float2(a.x * b.x, fmaf(a.x, b.y, a.y * b.x)); }
GPF_DEV du d_scale(float s, du a) { return make_float2(s * a.x, s * a.y); }
GPF_DEV du d_sp(du h) { return make_float2(fmaf(-h.x, h.x, 1.0f), -2.0f * h.x * h.y); }  // 1 - h^2
GPF_DEV du d_sq_minus(du x, float b) { return make_float2(fmaf(x.x, x.x, -b), 2.0f * x.x * x.y); }  // x^2 - b
#endif

template <int H1_, int H2_, int NW_>
struct SF2E {
  typedef SF2R<H1_, H2_, NW_> R;
  static constexpr int H1 = H1_, H2 = H2_, NW = NW_;
  // dual stash row: 16 chunks of 16 B; chunk c < 8: positions 2c, 2c+1 of (X0, h1[0..14]), chunk 8 + c: of (X1, h2[0..14])
  static constexpr int DROW = 64, D_FL = 32 * DROW, DSTASH_BYTES = D_FL * 4;   // 8 KB per tile and dual evaluation
  // stash regions: two of 8 KB per warp; a plain evaluation uses the first 4 KB of its region
  static constexpr int REG_FL = D_FL;
  // dual work row (92 floats = 23 x 16 B, odd): L (16 positions x pair) | C (16 x pair) | E (16) | XG (X0 X1 gb0 gb1 pairs)
  static constexpr int WS = 92, O_L = 0, O_C = 32, O_E = 64, O_XG = 80;
  static constexpr int WK_FL = 32 * WS;   // also holds the plain work rows (32 x 68)
  static_assert(WK_FL >= R::WK_FL, "plain work rows alias the dual ones");
  // NW > 8: ONE stash region per warp (no overlap of a bulk copy with the next evaluation of the same warp; more warps cover it)
  static constexpr bool SINGLE = NW_ > 8;
  static constexpr int NREG = SINGLE ? 1 : 2;
  static constexpr int WARP_FL = NREG * REG_FL + WK_FL;
  // stash slots of one RK4 step: 4 plain evaluations, then 4 dual ones (direction w: point 1, point 2; direction z: ...)
  static constexpr int STEP_BYTES = 4 * R::STASH_BYTES + 4 * DSTASH_BYTES;
  GPF_DEV static size_t op_offset(int step, int j) {
    return (size_t)step * STEP_BYTES + (j < 4 ? (size_t)j * R::STASH_BYTES : (size_t)4 * R::STASH_BYTES + (size_t)(j - 4) * DSTASH_BYTES);
  }
  GPF_DEV static uint32_t op_bytes(int j) { return j < 4 ? (uint32_t)R::STASH_BYTES : (uint32_t)DSTASH_BYTES; }
  static size_t smem_bytes(int n_steps) {
    return (size_t)(R::WEND + R::SCT + round_up(n_steps + 1, 4) + round_up(R::NPARAM, 4) + NW * WARP_FL) * sizeof(float) + 128;
  }
};

// swizzled float offset of chunk c (0..15) in a dual stash row
GPF_DEV int dsw_chunk(int c, int l7) { return ((c & 8) << 2) + (((c & 7) ^ l7) << 2); }

template <int BIT>
GPF_DEV void load_w16s(const float* W, int off, float (&w)[16]) {
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float4 t = ldwr4<BIT>(W, off + 4 * q);
    w[4 * q] = t.x; w[4 * q + 1] = t.y; w[4 * q + 2] = t.z; w[4 * q + 3] = t.w;
  }
}

// the closed form A.2 for d = 2, nb = 1 in dual arithmetic
GPF_DEV void sf2e_closed_form(du X0, du X1, du psi, du G0, du G1, float bnd, du& V0, du& V1) {
  const du s0 = d_sq_minus(X0, bnd), s1 = d_sq_minus(X1, bnd);
  const du q0 = d_add(d_scale(2.0f, d_mul(psi, X1)), d_mul(s1, G1));
  const du q1 = d_neg(d_add(d_scale(2.0f, d_mul(psi, X0)), d_mul(s0, G0)));
  V0 = d_mul(s0, q0);
  V1 = d_mul(s1, q1);
}

// ---------------------------------------------------------------------------------------------------------
// dual forward: V(X, ts), X = (x, direction).  Writes the dual stash row of this lane's sample.
// ---------------------------------------------------------------------------------------------------------
// Wrow: this lane's work row (free during the forward sweep): scratch for ws, so that the two contractions can be ROLLED
// loops over the contraction index (the fully unrolled evaluation was 1 500 instructions; with plain + dual, forward +
// reverse the kernel outgrew the instruction cache: no_instruction was the top stall, profiles/r2_sf2e_*).
template <class K>
GPF_DEV void sf2e_forward_eval(const float* __restrict__ W, float* __restrict__ Drow, float* __restrict__ Wrow, int l7, du X0,
                               du X1, float ts, float bnd, du& V0, du& V1) {
  typedef typename K::R R;
  constexpr int H1 = K::H1, H2 = K::H2;
  du hp[16];  // by position
  hp[0] = X0;
#pragma unroll
  for (int m = 0; m < 15; ++m) {
    if (m < H1) {
      const float4 w = ldwr4<1>(W, R::W1P + 4 * m);
      const du a = d_fma(w.x, X0, d_fma(w.y, X1, d_make(fmaf(w.z, ts, w.w), 0.0f)));
      const float h = tanh_real(a.x);
      hp[m + 1] = d_make(h, fmaf(-h, h, 1.0f) * a.y);
    } else {
      hp[m + 1] = d_make(0.0f, 0.0f);
    }
  }
#pragma unroll
  for (int c = 0; c < 8; ++c) sts4(Drow + dsw_chunk(c, l7), hp[2 * c].x, hp[2 * c].y, hp[2 * c + 1].x, hp[2 * c + 1].y);
  // ---- layer 2: a2[k] = b2[k] + sum_m W2[k][m] h1[m], one FFMA2 per (k, m): the pair is (value, epsilon)
  du a2[16];
  {
    float b2[16];
    load_w16s<1>(W, R::W2F, b2);
#pragma unroll
    for (int k = 0; k < 15; ++k) a2[k] = d_make(b2[k + 1], 0.0f);   // W2F columns are positions (unit + 1)
    a2[15] = d_make(0.0f, 0.0f);
  }
#pragma unroll kSF2EUnroll
  for (int p = 1; p <= H1; ++p) {
    float wr[16];
    load_w16s<GPF_SF2E_WBIT>(W, R::W2F + 16 * p, wr);
    const du h = *(reinterpret_cast<const float2*>(Drow + dsw_chunk(p >> 1, l7)) + (p & 1));   // hp[p], just stored
#pragma unroll
    for (int k = 0; k < H2; ++k) a2[k] = d_fma(wr[k + 1], h, a2[k]);
  }
  // ---- activation 2, collapsed output row
  du h2p[16], ws[16];  // by position
  h2p[0] = X1;
  ws[0] = d_make(0.0f, 0.0f);
  du psi = d_make(ldwr4<1>(W, R::B3C).x, 0.0f);
  {
    float w3[16];
    load_w16s<1>(W, R::W3P, w3);
#pragma unroll
    for (int p = 1; p < 16; ++p) {
      if (p - 1 < H2) {
        const float h = tanh_real(a2[p - 1].x);
        const du h2 = d_make(h, fmaf(-h, h, 1.0f) * a2[p - 1].y);
        psi = d_fma(w3[p], h2, psi);
        ws[p] = d_scale(w3[p], d_sp(h2));
        h2p[p] = h2;
      } else {
        h2p[p] = ws[p] = d_make(0.0f, 0.0f);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    sts4(Drow + dsw_chunk(8 + c, l7), h2p[2 * c].x, h2p[2 * c].y, h2p[2 * c + 1].x, h2p[2 * c + 1].y);
    sts4(Wrow + 4 * c, ws[2 * c].x, ws[2 * c].y, ws[2 * c + 1].x, ws[2 * c + 1].y);
  }
  // ---- z = W2^T ws, by position of h1
  du z[16];
#pragma unroll
  for (int pm = 0; pm < 16; ++pm) z[pm] = d_make(0.0f, 0.0f);
#pragma unroll kSF2EUnroll
  for (int p = 1; p <= H2; ++p) {
    float wr[16];
    load_w16s<GPF_SF2E_WBIT>(W, R::W2B + 16 * p, wr);
    const du wsp = *reinterpret_cast<const float2*>(Wrow + 2 * p);
#pragma unroll
    for (int pm = 1; pm <= H1; ++pm) z[pm] = d_fma(wr[pm], wsp, z[pm]);
  }
  // ---- grad psi = W1x^T (sp1 (.) z)
  du G0 = d_make(0.0f, 0.0f), G1 = d_make(0.0f, 0.0f);
#pragma unroll
  for (int m = 0; m < H1; ++m) {
    const float4 w = ldwr4<1>(W, R::W1P + 4 * m);
    const du q = d_mul(d_sp(hp[m + 1]), z[m + 1]);
    G0 = d_fma(w.x, q, G0);
    G1 = d_fma(w.y, q, G1);
  }
  sf2e_closed_form(X0, X1, psi, G0, G1, bnd, V0, V1);
}

}  // namespace gpflow
