// gpflow_b200 -- sf2r kernel, reverse sweep of one velocity evaluation for one warp tile (formulas: gpflow_sf2r.cuh).
#pragma once
#include "gpflow_generic_mma.cuh"   // mma_tf32 (m16n8k8 TF32 fragments)
#include "gpflow_sf2r.cuh"

namespace gpflow {

// Given the stash rows of one evaluation and the cotangent (vb0, vb1) of v, returns xbar = J_v^T vbar and accumulates
// every parameter gradient:
//   acc[16]   lane-owned 4x4 tile (positions 4ti.., 4tj..) of dW2 (+ db2 in column 0); rank-64 update over the tile
//   accw[16]  lane-owned 4x4 tile {dW1[m][0], dW1[m][1], dW1[m][2] (time), db1[m]}, POSITION m + 1 = 4 (lane & 3) + r
//   acc3[4]   dw3 of the collapsed output row at positions 4 (lane & 3) + r (dW3[0] = +, dW3[1] = -);  b3b lane-private
template <class K>
GPF_DEV void sf2r_reverse_eval(const float* __restrict__ W, const float* __restrict__ Sbuf, float* __restrict__ Wk, int lane,
                               float ts, float bnd, float vb0, float vb1, float& xbar0, float& xbar1, float (&acc)[16],
                               float (&accw)[16], float (&acc3)[4], float& b3b) {
  constexpr int H1 = K::H1;
  const int l7 = lane & 7;
  const float* Srow = Sbuf + lane * K::SROW;
  float* Wrow = Wk + lane * K::WS;
  float2 hpp[8];  // position pairs: hpp[0].x = x0, then h1
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const float4 t = lds4(Srow + sw_chunk(c, l7));
    hpp[2 * c] = make_float2(t.x, t.y);
    hpp[2 * c + 1] = make_float2(t.z, t.w);
  }
  const float x0 = hpp[0].x, x1 = Srow[sw_chunk(4, l7)];
  // ---- A.3, the part that does not need psi / grad psi
  const float s0 = fmaf(x0, x0, -bnd), s1 = fmaf(x1, x1, -bnd);
  const float qb0 = vb0 * s0, qb1 = vb1 * s1;
  const float psib = 2.0f * (x1 * qb0 - x0 * qb1);
  const float gb1 = s1 * qb0, gb0 = -s0 * qb1;
  b3b += psib;
  // ---- c = sp1 (.) (W1x gb) as position pairs (the layer-1 weights of position 0 are zero, so c[0] = 0): the pairs are those
  //      of the 128-bit stores and of the position-indexed weight rows; the work row also gets (x0 x1 gb0 gb1).  Position 0
  //      of this sample's stash row (x0, already read) becomes 1.0: the bias column of the dW2 update below
  const float2 one2 = make_float2(1.0f, 1.0f), gb0d = make_float2(gb0, gb0), gb1d = make_float2(gb1, gb1);
  float2 cp[8], sp1[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float4 w0 = ldwr4<4>(W, K::W1Q + 8 * j);
    sp1[j] = __ffma2_rn(make_float2(-hpp[j].x, -hpp[j].y), hpp[j], one2);
    cp[j] = __fmul2_rn(sp1[j], __ffma2_rn(make_float2(w0.x, w0.y), gb0d, __fmul2_rn(make_float2(w0.z, w0.w), gb1d)));
  }
#pragma unroll
  for (int c = 0; c < 4; ++c) sts4(Wrow + K::O_C + 4 * c, cp[2 * c].x, cp[2 * c].y, cp[2 * c + 1].x, cp[2 * c + 1].y);
  sts4(Wrow + K::O_XG, x0, x1, gb0, gb1);
  const_cast<float*>(Srow)[sw_chunk(0, l7)] = 1.0f;
  // ---- one pass over the rows of W2 (position p of h2, rolled by chunk): tb = W2 c (dot), then hb += W2[k,:] ab2[k],
  //      z += W2[k,:] ws[k].  Position 0 (the x1 slot) has a zero weight row and w3 = 0: it contributes nothing.
  float2 hbp[8], zp[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) hbp[j] = zp[j] = make_float2(0.0f, 0.0f);
  float psi = ldwr4<4>(W, K::B3C).x;
  GPF_SF2_LOOP
  for (int c = 0; c < 4; ++c) {
    const float4 hc = lds4(Srow + sw_chunk(4 + c, l7));
    const float4 w3v = ldwr4<4>(W, K::W3P + 4 * c);
    const float hh[4] = {hc.x, hc.y, hc.z, hc.w}, w3[4] = {w3v.x, w3v.y, w3v.z, w3v.w};
    float ab[4], wsv[4], ev[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (i == 0 && c == 0) {   // position 0 is the x1 slot (zero weight row, w3 = 0): skip it (warp-uniform branch)
        ab[0] = wsv[0] = ev[0] = 0.0f;
        continue;
      }
      float2 wp[8];
      load_w16<2>(W, K::W2B + 16 * (4 * c + i), wp);
      float2 ta = __fmul2_rn(wp[0], cp[0]), tc = __fmul2_rn(wp[1], cp[1]);
#pragma unroll
      for (int j = 1; j < 4; ++j) {
        ta = __ffma2_rn(wp[2 * j], cp[2 * j], ta);
        tc = __ffma2_rn(wp[2 * j + 1], cp[2 * j + 1], tc);
      }
      const float tb = (ta.x + tc.x) + (ta.y + tc.y);
      const float h = hh[i], sp = fmaf(-h, h, 1.0f);
      ev[i] = fmaf(psib, h, sp * tb);                    // dw3 contribution: psib h2 + sp2 tb
      const float wsi = w3[i] * sp;
      const float abi = wsi * fmaf(-2.0f * h, tb, psib);   // ab2 = sp2 (psib w3 - 2 h2 w3 tb)
      psi = fmaf(w3[i], h, psi);
      const float2 abd = make_float2(abi, abi), wsd = make_float2(wsi, wsi);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        hbp[j] = __ffma2_rn(wp[j], abd, hbp[j]);
        zp[j] = __ffma2_rn(wp[j], wsd, zp[j]);
      }
      ab[i] = abi;
      wsv[i] = wsi;
    }
    sts4(Wrow + K::O_AB + 4 * c, ab[0], ab[1], ab[2], ab[3]);
    sts4(Wrow + K::O_WS + 4 * c, wsv[0], wsv[1], wsv[2], wsv[3]);
    sts4(Wrow + K::O_E + 4 * c, ev[0], ev[1], ev[2], ev[3]);
  }
  // ---- activation 1 + layer 1 as position pairs: q = sp1 z, grad psi, ab1 = sp1 hb - 2 h1 z c, xbar
  //      (position 0: z = hb = c = 0 and zero layer-1 weights, so its ab1 = q = 0)
  float2 g0p = make_float2(0.0f, 0.0f), g1p = g0p, xb0p = g0p, xb1p = g0p;
  float2 ab1[8], qv[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float4 w0 = ldwr4<4>(W, K::W1Q + 8 * j);
    const float2 wa = make_float2(w0.x, w0.y), wb = make_float2(w0.z, w0.w);
    const float2 q = __fmul2_rn(sp1[j], zp[j]);
    g0p = __ffma2_rn(wa, q, g0p);
    g1p = __ffma2_rn(wb, q, g1p);
    const float2 hz = __fmul2_rn(__fmul2_rn(hpp[j], zp[j]), make_float2(-2.0f, -2.0f));
    const float2 a = __ffma2_rn(hz, cp[j], __fmul2_rn(sp1[j], hbp[j]));
    xb0p = __ffma2_rn(wa, a, xb0p);
    xb1p = __ffma2_rn(wb, a, xb1p);
    ab1[j] = a;
    qv[j] = q;
  }
  const float g0 = g0p.x + g0p.y, g1 = g1p.x + g1p.y;
  float xb0 = xb0p.x + xb0p.y, xb1 = xb1p.x + xb1p.y;
  // ---- rest of A.3
  {
    const float q0 = fmaf(2.0f * psi, x1, s1 * g1);
    const float q1 = -fmaf(2.0f * psi, x0, s0 * g0);
    const float sb0 = fmaf(vb0, q0, -g0 * qb1), sb1 = fmaf(vb1, q1, g1 * qb0);
    xb0 += fmaf(2.0f * x0, sb0, -2.0f * psi * qb1);
    xb1 += fmaf(2.0f * x1, sb1, 2.0f * psi * qb0);
  }
  __syncwarp();
#if GPF_SF2R_MMA
  // ---- dW2[pk][pm] += sum_s ab2_s[pk] h1_s[pm] + ws_s[pk] c_s[pm]  (h1 position 0 == 1 -> db2) on the tensor cores:
  //      a 16 x 16 x 64 product (positions x positions x (32 samples, 2 terms)) as mma.sync m16n8k8 TF32 tiles with the
  //      3xTF32 split (fp32-level accuracy, gpflow_generic_mma.cuh).  It takes the rank update (19 % of the FMA-pipe
  //      cycles, 30 % of the shared-memory wavefronts) off the two pipes that bound the kernel.  Contraction index c of
  //      k-step ks <-> sample 8 ks + 2 (c mod 4) + (c div 4): with that order every fragment load is a conflict-free
  //      32-bit shared-memory access (work rows: stride 68 floats; stash rows: the XOR swizzle).  The products of ONE
  //      evaluation are accumulated from zero in the fragment and then added in fp32 (no long accumulation chains inside
  //      the tensor core).  acc[0..7] = the m16n8 C fragments of the two column tiles: lane (g, t) owns the entries
  //      (pk in {g, g + 8}, pm in {2t, 2t + 1} + 8 nt).
  {
    const int g = lane >> 2, t = lane & 3;
    float cf[4][2][4];   // four independent accumulation chains (k-step parity x term): mma latency, not throughput, binds
#pragma unroll
    for (int ch = 0; ch < 4; ++ch)
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int i = 0; i < 4; ++i) cf[ch][nt][i] = 0.0f;
    auto split = [](float x, uint32_t& hi, uint32_t& lo) {
#if GPF_SF2R_MMA == 2
      hi = __float_as_uint(x) & 0xffffe000u;             // truncation split (ALU) instead of cvt.rna
#else
      asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(x));
#endif
      lo = __float_as_uint(x - __uint_as_float(hi));   // the tensor core reads the top 19 bits: 2^-21 relative at worst
    };
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      const int sa = 8 * ks + 2 * t, sb = sa + 1;
      const float* wa = Wk + sa * K::WS;
      const float* wb = Wk + sb * K::WS;
      const float* da = Sbuf + sa * K::SROW;
      const float* db = Sbuf + sb * K::SROW;
#pragma unroll
      for (int term = 0; term < 2; ++term) {
        const int oa = term == 0 ? K::O_AB : K::O_WS;
        uint32_t ah[4], al[4];
        split(wa[oa + g], ah[0], al[0]);
        split(wa[oa + g + 8], ah[1], al[1]);
        split(wb[oa + g], ah[2], al[2]);
        split(wb[oa + g + 8], ah[3], al[3]);
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          const int pm = 8 * nt + g;
          float b0, b1;
          if (term == 0) {
            b0 = da[sw_chunk(pm >> 2, sa & 7) + (pm & 3)];
            b1 = db[sw_chunk(pm >> 2, sb & 7) + (pm & 3)];
          } else {
            b0 = wa[K::O_C + pm];
            b1 = wb[K::O_C + pm];
          }
          uint32_t bh[2], bl[2];
          split(b0, bh[0], bl[0]);
          split(b1, bh[1], bl[1]);
          mma_3xtf32(cf[2 * term + (ks & 1)][nt], ah, al, bh, bl);
        }
      }
    }
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[4 * nt + i] += (cf[0][nt][i] + cf[1][nt][i]) + (cf[2][nt][i] + cf[3][nt][i]);
  }
#else
  // ---- dW2[pk][pm] += sum_s ab2_s[pk] h1_s[pm] + ws_s[pk] c_s[pm]  (h1 position 0 == 1 -> db2): lane (slice, ti, tj)
  //      owns the 4x4 tile (4ti.., 4tj..) over the 16 samples with bit 2 == slice (conflict-free with the swizzle)
  {
    const int tj = lane & 3, ti = (lane >> 2) & 3, slice = lane >> 4;
    const float* wl = Wk + 4 * slice * K::WS + 4 * ti;                 // left factors:  + O_AB / O_WS
    const float* wr = Wk + 4 * slice * K::WS + K::O_C + 4 * tj;        // right factor c
    const float* sr[4];                                                // right factor h1: sample 4 slice + ql (+ 8 qh)
#pragma unroll
    for (int ql = 0; ql < 4; ++ql) sr[ql] = Sbuf + (4 * slice + ql) * K::SROW + sw_chunk(tj, 4 * slice + ql);
    float2 accp[4][2];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      accp[r][0] = make_float2(acc[4 * r], acc[4 * r + 1]);
      accp[r][1] = make_float2(acc[4 * r + 2], acc[4 * r + 3]);
    }
#pragma unroll
    for (int qh = 0; qh < 4; ++qh) {
#pragma unroll
      for (int ql = 0; ql < 4; ++ql) {
        const int so = ql + 8 * qh;
        const float4 rh = lds4(sr[ql] + 8 * qh * K::SROW), rc = lds4(wr + so * K::WS);
        const float4 la = lds4(wl + so * K::WS + K::O_AB), lw = lds4(wl + so * K::WS + K::O_WS);
        const float2 r0[2] = {make_float2(rh.x, rh.y), make_float2(rh.z, rh.w)};
        const float2 r1[2] = {make_float2(rc.x, rc.y), make_float2(rc.z, rc.w)};
        const float a4[4] = {la.x, la.y, la.z, la.w}, w4[4] = {lw.x, lw.y, lw.z, lw.w};
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const float2 ad = make_float2(a4[r], a4[r]), wd = make_float2(w4[r], w4[r]);
#pragma unroll
          for (int i = 0; i < 2; ++i) accp[r][i] = __ffma2_rn(ad, r0[i], __ffma2_rn(wd, r1[i], accp[r][i]));
        }
      }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      acc[4 * r] = accp[r][0].x; acc[4 * r + 1] = accp[r][0].y;
      acc[4 * r + 2] = accp[r][1].x; acc[4 * r + 3] = accp[r][1].y;
    }
  }
#endif
  __syncwarp();  // every lane is done with the other lanes' [ab2 | ws] rows
#pragma unroll
  for (int c = 0; c < 4; ++c) {   // by position (row 0 of the dW1 tile is dead)
    sts4(Wrow + K::O_AB + 4 * c, ab1[2 * c].x, ab1[2 * c].y, ab1[2 * c + 1].x, ab1[2 * c + 1].y);
    sts4(Wrow + K::O_WS + 4 * c, qv[2 * c].x, qv[2 * c].y, qv[2 * c + 1].x, qv[2 * c + 1].y);
  }
  __syncwarp();
  // ---- dW1[m][:] += sum_s ab1_s[m] [x0, x1, t, 1]_s + q_s[m] [gb0, gb1, 0, 0]_s;  dw3[p] += sum_s e_s[p]:
  //      lane (slice, ti) owns positions 4ti.. (unit + 1) over the samples s = slice (mod 8)
  {
    const int ti = lane & 3, slice = lane >> 2;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float* wr = Wk + (slice + 8 * q) * K::WS;
      const float4 xg = lds4(wr + K::O_XG);
      const float4 l0 = lds4(wr + K::O_AB + 4 * ti), l1 = lds4(wr + K::O_WS + 4 * ti), l2 = lds4(wr + K::O_E + 4 * ti);
      const float la[4] = {l0.x, l0.y, l0.z, l0.w}, lq[4] = {l1.x, l1.y, l1.z, l1.w}, le[4] = {l2.x, l2.y, l2.z, l2.w};
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        accw[4 * r + 0] = fmaf(la[r], xg.x, fmaf(lq[r], xg.z, accw[4 * r + 0]));
        accw[4 * r + 1] = fmaf(la[r], xg.y, fmaf(lq[r], xg.w, accw[4 * r + 1]));
        accw[4 * r + 2] = fmaf(la[r], ts, accw[4 * r + 2]);
        accw[4 * r + 3] += la[r];
        acc3[r] += le[r];
      }
    }
  }
  xbar0 = xb0;
  xbar1 = xb1;
}

}  // namespace gpflow
