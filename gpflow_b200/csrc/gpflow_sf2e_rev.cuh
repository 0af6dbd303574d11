// gpflow_b200 -- sf2e kernel: reverse sweep of one DUAL velocity evaluation for one warp tile (see gpflow_sf2e.cuh).
// sf2r_reverse_eval with every product replaced by the dual product; the parameter gradients that are accumulated are the
// EPSILON parts of the dual gradients (the gradient of the directional derivative).
#pragma once
#include "gpflow_sf2e.cuh"

namespace gpflow {

// Dbuf: dual stash rows of the evaluation; Wk: dual work rows [32][K::WS]; (Vb0, Vb1): dual cotangent of V.
// Returns the dual cotangent of X and adds the epsilon parts of the parameter gradients to the accumulators of
// sf2r_reverse_eval (acc: dW2 tile by position, accw: dW1 tile by unit, acc3: dw3 by position, b3b).
template <class K>
GPF_DEV void sf2e_reverse_eval(const float* __restrict__ W, float* __restrict__ Dbuf, float* __restrict__ Wk, int lane, float ts,
                               float bnd, du Vb0, du Vb1, du& Xbar0, du& Xbar1, float (&acc)[16], float (&accw)[16],
                               float (&acc3)[4], float& b3b) {
  typedef typename K::R R;
  constexpr int H1 = K::H1, H2 = K::H2;
  const int l7 = lane & 7;
  float* Drow = Dbuf + lane * K::DROW;
  float* Wrow = Wk + lane * K::WS;
  du hp[16];  // by position: X0, h1[0..14]
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const float4 t = lds4(Drow + dsw_chunk(c, l7));
    hp[2 * c] = d_make(t.x, t.y);
    hp[2 * c + 1] = d_make(t.z, t.w);
  }
  const du X0 = hp[0];
  float2* const B0 = reinterpret_cast<float2*>(Drow + dsw_chunk(8, l7));   // position 0 of the (X1, h2) half
  const du X1 = *B0;
  // ---- A.3 in dual arithmetic, the part that does not need psi / grad psi
  const du s0 = d_sq_minus(X0, bnd), s1 = d_sq_minus(X1, bnd);
  const du qb0 = d_mul(Vb0, s0), qb1 = d_mul(Vb1, s1);
  const du psib = d_scale(2.0f, d_sub(d_mul(X1, qb0), d_mul(X0, qb1)));
  const du gb1 = d_mul(s1, qb0), gb0 = d_neg(d_mul(s0, qb1));
  b3b += psib.y;
  // ---- c = sp1 (.) (W1x gb) by position (c[0] = 0); C row stores it as (e, v): right factor of the second dW2 term
  du cq[16];
  cq[0] = d_make(0.0f, 0.0f);
#pragma unroll
  for (int m = 0; m < 15; ++m) {
    if (m < H1) {
      const float4 w = ldwr4<GPF_SF2E_W1BIT>(W, R::W1P + 4 * m);
      cq[m + 1] = d_mul(d_sp(hp[m + 1]), d_fma(w.x, gb0, d_scale(w.y, gb1)));
    } else {
      cq[m + 1] = d_make(0.0f, 0.0f);
    }
  }
#pragma unroll
  for (int c = 0; c < 8; ++c) sts4(Wrow + K::O_C + 4 * c, cq[2 * c].y, cq[2 * c].x, cq[2 * c + 1].y, cq[2 * c + 1].x);
  sts4(Wrow + K::O_XG, X0.x, X0.y, X1.x, X1.y);
  sts4(Wrow + K::O_XG + 4, gb0.x, gb0.y, gb1.x, gb1.y);
  // position 0 of the stash row (X0, already read) becomes the dual number 1: the db2 column of the dW2 update
  *reinterpret_cast<float2*>(Drow + dsw_chunk(0, l7)) = make_float2(1.0f, 0.0f);
  *B0 = make_float2(0.0f, 0.0f);   // ... and the (X1, h2) half is rewritten in place into ws (left factor of the second term)
  // ---- one pass over the rows of W2 (position p of h2): tb = W2 c, ab2, ws, then hb += W2[k,:] ab2[k], z += W2[k,:] ws[k]
  du hb[16], z[16];
#pragma unroll
  for (int pm = 0; pm < 16; ++pm) hb[pm] = z[pm] = d_make(0.0f, 0.0f);
  du psi = d_make(ldwr4<GPF_SF2E_W1BIT>(W, R::B3C).x, 0.0f);
  Wrow[K::O_E] = 0.0f;
  sts4(Wrow + K::O_L, 0.0f, 0.0f, 0.0f, 0.0f);   // position 0 of the left factor
  for (int p = H2 + 1; p < 16; ++p) {            // positions beyond the last unit (none for H2 = 15)
    Wrow[K::O_E + p] = 0.0f;
    *reinterpret_cast<float2*>(Wrow + K::O_L + 2 * p) = make_float2(0.0f, 0.0f);
  }
#pragma unroll kSF2EUnroll
  for (int p = 1; p <= H2; ++p) {   // ROLLED over the rows of W2 (code size, see sf2e_forward_eval)
    float wr[16];
    load_w16s<GPF_SF2E_RBIT>(W, R::W2B + 16 * p, wr);
    du ta = d_scale(wr[1], cq[1]), tc = d_scale(wr[2], cq[2]), td = d_scale(wr[3], cq[3]);
#pragma unroll
    for (int j = 1; j < 5; ++j) {
      ta = d_fma(wr[3 * j + 1], cq[3 * j + 1], ta);
      tc = d_fma(wr[3 * j + 2], cq[3 * j + 2], tc);
      td = d_fma(wr[3 * j + 3], cq[3 * j + 3], td);
    }
    const du tb = d_add(d_add(ta, tc), td);
    const float w3p = ldsw1(W + R::W3P + p);
    float2* const hslot = reinterpret_cast<float2*>(Drow + dsw_chunk(8 + (p >> 1), l7)) + (p & 1);
    const du h2 = *hslot, sp = d_sp(h2);
    Wrow[K::O_E + p] = fmaf(psib.x, h2.y, fmaf(psib.y, h2.x, fmaf(sp.x, tb.y, sp.y * tb.x)));   // eps part of psib h2 + sp2 tb
    const du wsp = d_scale(w3p, sp);
    const du ab = d_mul(wsp, d_sub(psib, d_scale(2.0f, d_mul(h2, tb))));
    psi = d_fma(w3p, h2, psi);
#pragma unroll
    for (int pm = 1; pm <= H1; ++pm) {
      hb[pm] = d_fma(wr[pm], ab, hb[pm]);
      z[pm] = d_fma(wr[pm], wsp, z[pm]);
    }
    *hslot = wsp;                                                                   // left factor of term 2: (v, e)
    *reinterpret_cast<float2*>(Wrow + K::O_L + 2 * p) = make_float2(ab.y, ab.x);   // left factor of term 1: (e, v)
  }
  // ---- eps part of dW2[pk][pm] = sum_s ab2.v h1.e + ab2.e h1.v + ws.v c.e + ws.e c.v: a rank-64 update over the 16 samples
  //      with bit 0 == slice; the pair accumulators hold the two halves of each entry
  {
    const int tj = lane & 3, ti = (lane >> 2) & 3, slice = lane >> 4;
    float2 accp[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) accp[r][c] = make_float2(0.0f, 0.0f);
    const float* wl = Wk + slice * K::WS + K::O_L + 8 * ti;
    const float* wc = Wk + slice * K::WS + K::O_C + 8 * tj;
    __syncwarp();
#pragma unroll kSF2ERankUnroll
    for (int q = 0; q < 16; ++q) {
      const int s = slice + 2 * q;
      const float* dr = Dbuf + s * K::DROW;
      {   // term 1: (ab2.e, ab2.v) x (h1.v, h1.e)
        const float4 l0 = lds4(wl + 2 * q * K::WS), l1 = lds4(wl + 2 * q * K::WS + 4);
        const float4 r0 = lds4(dr + dsw_chunk(2 * tj, s & 7)), r1 = lds4(dr + dsw_chunk(2 * tj + 1, s & 7));
        const float2 L[4] = {make_float2(l0.x, l0.y), make_float2(l0.z, l0.w), make_float2(l1.x, l1.y), make_float2(l1.z, l1.w)};
        const float2 Rr[4] = {make_float2(r0.x, r0.y), make_float2(r0.z, r0.w), make_float2(r1.x, r1.y), make_float2(r1.z, r1.w)};
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
          for (int c = 0; c < 4; ++c) accp[r][c] = __ffma2_rn(L[r], Rr[c], accp[r][c]);
      }
      {   // term 2: (ws.v, ws.e) x (c.e, c.v)
        const float4 l0 = lds4(dr + dsw_chunk(8 + 2 * ti, s & 7)), l1 = lds4(dr + dsw_chunk(9 + 2 * ti, s & 7));
        const float4 r0 = lds4(wc + 2 * q * K::WS), r1 = lds4(wc + 2 * q * K::WS + 4);
        const float2 L[4] = {make_float2(l0.x, l0.y), make_float2(l0.z, l0.w), make_float2(l1.x, l1.y), make_float2(l1.z, l1.w)};
        const float2 Rr[4] = {make_float2(r0.x, r0.y), make_float2(r0.z, r0.w), make_float2(r1.x, r1.y), make_float2(r1.z, r1.w)};
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
          for (int c = 0; c < 4; ++c) accp[r][c] = __ffma2_rn(L[r], Rr[c], accp[r][c]);
      }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[4 * r + c] += accp[r][c].x + accp[r][c].y;
  }
  __syncwarp();  // every lane is done with the other lanes' ab2 / c rows: they become ab1 / q (by position, (v, e))
  // ---- activation 1 + layer 1: q = sp1 z, grad psi, ab1, xbar
  du G0 = d_make(0.0f, 0.0f), G1 = G0, Xb0 = G0, Xb1 = G0;
  sts4(Wrow + K::O_L, 0.0f, 0.0f, 0.0f, 0.0f);   // position 0 of ab1 / q (the rows of the dW1 tile are positions)
  sts4(Wrow + K::O_C, 0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll
  for (int c = 0; c < 8; ++c) {   // h1 again (position 0 is the constant 1 now; X0 is in a register)
    const float4 t = lds4(Drow + dsw_chunk(c, l7));
    hp[2 * c] = d_make(t.x, t.y);
    hp[2 * c + 1] = d_make(t.z, t.w);
  }
#pragma unroll
  for (int m = 0; m < 16; ++m) {
    if (m < H1) {
      const float4 w = ldwr4<GPF_SF2E_W1BIT>(W, R::W1P + 4 * m);
      const du h = hp[m + 1], sp = d_sp(h);
      const du q = d_mul(sp, z[m + 1]);
      G0 = d_fma(w.x, q, G0);
      G1 = d_fma(w.y, q, G1);
      const du a = d_sub(d_mul(sp, hb[m + 1]), d_scale(2.0f, d_mul(d_mul(h, z[m + 1]), cq[m + 1])));
      Xb0 = d_fma(w.x, a, Xb0);
      Xb1 = d_fma(w.y, a, Xb1);
      *reinterpret_cast<float2*>(Wrow + K::O_L + 2 * (m + 1)) = a;   // by position, (v, e)
      *reinterpret_cast<float2*>(Wrow + K::O_C + 2 * (m + 1)) = q;
    } else if (m < 15) {
      *reinterpret_cast<float2*>(Wrow + K::O_L + 2 * (m + 1)) = make_float2(0.0f, 0.0f);
      *reinterpret_cast<float2*>(Wrow + K::O_C + 2 * (m + 1)) = make_float2(0.0f, 0.0f);
    }
  }
  // ---- rest of A.3
  {
    const du q0 = d_add(d_scale(2.0f, d_mul(psi, X1)), d_mul(s1, G1));
    const du q1 = d_neg(d_add(d_scale(2.0f, d_mul(psi, X0)), d_mul(s0, G0)));
    const du sb0 = d_sub(d_mul(Vb0, q0), d_mul(G0, qb1)), sb1 = d_add(d_mul(Vb1, q1), d_mul(G1, qb0));
    Xb0 = d_add(Xb0, d_scale(2.0f, d_sub(d_mul(X0, sb0), d_mul(psi, qb1))));
    Xb1 = d_add(Xb1, d_scale(2.0f, d_add(d_mul(X1, sb1), d_mul(psi, qb0))));
  }
  __syncwarp();
  // ---- eps part of dW1[m][j] = sum_s (ab1 (x) X_j + q (x) gb_j).e, time column ab1.e t, db1 = ab1.e; dw3[p] += e[p]
  {
    const int ti = lane & 3, slice = lane >> 2;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float* wr = Wk + (slice + 8 * q) * K::WS;
      const float4 xx = lds4(wr + K::O_XG), gg = lds4(wr + K::O_XG + 4);   // X0.v X0.e X1.v X1.e | gb0.v gb0.e gb1.v gb1.e
      const float4 a0 = lds4(wr + K::O_L + 8 * ti), a1 = lds4(wr + K::O_L + 8 * ti + 4);
      const float4 q0 = lds4(wr + K::O_C + 8 * ti), q1 = lds4(wr + K::O_C + 8 * ti + 4);
      const float4 e4 = lds4(wr + K::O_E + 4 * ti);
      const float av[4] = {a0.x, a0.z, a1.x, a1.z}, ae[4] = {a0.y, a0.w, a1.y, a1.w};
      const float qvv[4] = {q0.x, q0.z, q1.x, q1.z}, qe[4] = {q0.y, q0.w, q1.y, q1.w};
      const float le[4] = {e4.x, e4.y, e4.z, e4.w};
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        accw[4 * r + 0] = fmaf(av[r], xx.y, fmaf(ae[r], xx.x, fmaf(qvv[r], gg.y, fmaf(qe[r], gg.x, accw[4 * r + 0]))));
        accw[4 * r + 1] = fmaf(av[r], xx.w, fmaf(ae[r], xx.z, fmaf(qvv[r], gg.w, fmaf(qe[r], gg.z, accw[4 * r + 1]))));
        accw[4 * r + 2] = fmaf(ae[r], ts, accw[4 * r + 2]);
        accw[4 * r + 3] += ae[r];
        acc3[r] += le[r];
      }
    }
  }
  Xbar0 = Xb0;
  Xbar1 = Xb1;
}

}  // namespace gpflow
