// gpflow_b200 -- sf2 kernel, per-evaluation device code: velocity forward (A.1 + A.2) and its reverse sweep
// (A.3 + A.4) for one warp tile.  See gpflow_sf2.cuh for the layout.
#pragma once
#include "gpflow_sf2.cuh"

namespace gpflow {

// ---------------------------------------------------------------------------------------------------------
// forward: v(x, ts).  Writes h1 into the A row (always: layer 2 reads it back with a dynamic index) and, when
// STASH, the rest of the A/B rows the reverse sweep needs.
// ---------------------------------------------------------------------------------------------------------
template <class K, bool STASH>
GPF_DEV void sf2_forward_eval(const float* __restrict__ W, float* __restrict__ Arow, float* __restrict__ Brow, float x0,
                              float x1, float ts, float bnd, float& v0, float& v1) {
  constexpr int H1 = K::H1, H2 = K::H2;
  // ---- layer 1, output unit k unrolled: a1 = W1 [x; t] + b1, h1 = tanh(a1)   (vector_field.py:75-90)
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    float hv[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int k = 4 * c + i;
      if (k < H1) {
        const float4 w = ldw4(W, K::W1P + 4 * k);
        hv[i] = tanh_real(fmaf(w.x, x0, fmaf(w.y, x1, fmaf(w.z, ts, w.w))));
      } else {
        hv[i] = (k == 15) ? 1.0f : 0.0f;  // slot 15 == 1: bias column of layer 2 and of its weight gradient
      }
    }
    sts4(Arow + 4 + 4 * c, hv[0], hv[1], hv[2], hv[3]);
  }
  // ---- layer 2, rolled over the input unit m: a2 += W2[:,m] h1[m], U2[:,j] += W2[:,m] T1[m][j], T1 = sp1 W1x
  //      Packed fp32x2 FMAs (fma.rn.f32x2, SASS FFMA2 -- new on sm_100): one instruction = two output units.
  float2 a2p[8], u0p[8], u1p[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) a2p[j] = u0p[j] = u1p[j] = make_float2(0.0f, 0.0f);
  GPF_SF2_LOOP
  for (int mc = 0; mc < 4; ++mc) {
    const float4 hc = lds4(Arow + 4 + 4 * mc);
    const float hh[4] = {hc.x, hc.y, hc.z, hc.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int m = 4 * mc + i;
      const float4 w1 = ldw4(W, K::W1P + 4 * m);
      const float h = hh[i], sp = fmaf(-h, h, 1.0f), in1 = sp * w1.x, in2 = sp * w1.y;
      const float2 hd = make_float2(h, h), d1 = make_float2(in1, in1), d2 = make_float2(in2, in2);
      float2 wp[8];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 t = ldw4(W, K::W2T + 16 * m + 4 * q);
        wp[2 * q] = make_float2(t.x, t.y);
        wp[2 * q + 1] = make_float2(t.z, t.w);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        a2p[j] = __ffma2_rn(wp[j], hd, a2p[j]);
        u0p[j] = __ffma2_rn(wp[j], d1, u0p[j]);
        u1p[j] = __ffma2_rn(wp[j], d2, u1p[j]);
      }
    }
  }
  float a2[16], u0[16], u1[16];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    a2[2 * j] = a2p[j].x; a2[2 * j + 1] = a2p[j].y;
    u0[2 * j] = u0p[j].x; u0[2 * j + 1] = u0p[j].y;
    u1[2 * j] = u1p[j].x; u1[2 * j + 1] = u1p[j].y;
  }
  // ---- activation 2 + collapsed output layer: psi = w3 . h2 + b3c, g_j = w3 . (sp2 U2[:,j])
  float psi = ldw4(W, K::B3C).x, g0 = 0.0f, g1 = 0.0f;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const float4 w3v = ldw4(W, K::W3C + 4 * c);
    const float w3[4] = {w3v.x, w3v.y, w3v.z, w3v.w};
    float hv[4], p0[4], p1[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int k = 4 * c + i;
      if (k < H2) {
        const float h = tanh_real(a2[k]);
        const float sp = fmaf(-h, h, 1.0f);
        psi = fmaf(w3[i], h, psi);
        g0 = fmaf(w3[i], sp * u0[k], g0);
        g1 = fmaf(w3[i], sp * u1[k], g1);
        hv[i] = h; p0[i] = u0[k]; p1[i] = u1[k];
      } else {
        hv[i] = p0[i] = p1[i] = 0.0f;
      }
    }
    if (STASH) {
      if (c == 3) hv[3] = g1;  // B slot 15 carries d_1 psi
      sts4(Brow + 4 * c, hv[0], hv[1], hv[2], hv[3]);
      sts4(Brow + 16 + 4 * c, p0[0], p0[1], p0[2], p0[3]);
      sts4(Brow + 32 + 4 * c, p1[0], p1[1], p1[2], p1[3]);
    }
  }
  if (STASH) sts4(Arow, x0, x1, psi, g0);
  // ---- closed form A.2 for d = 2, nb = 1 (velocity_field.py:20-79)
  const float s0 = fmaf(x0, x0, -bnd), s1 = fmaf(x1, x1, -bnd);
  const float q0 = fmaf(2.0f * psi, x1, s1 * g1);
  const float q1 = -fmaf(2.0f * psi, x0, s0 * g0);
  v0 = s0 * q0;
  v1 = s1 * q1;
}

// ---------------------------------------------------------------------------------------------------------
// reverse: given the stashed rows of one evaluation (A: x0 x1 psi g0 | h1, B: h2,g1 | U2) and the cotangent
// (vb0, vb1) of v, returns xbar = J_v^T vbar and accumulates every parameter gradient.
//   acc[16]  lane-owned 4x4 tile of dW2 (+ db2 in column 15), rank-96 update over the tile's samples
//   w3b/b3b  lane-private sums for the collapsed output row (dW3[0] = +, dW3[1] = -)
//   accw[16] lane-owned 4x4 tile {dW1[m][0], dW1[m][1], dW1[m][2] (time), db1[m]}, m = 4 (lane & 3) + r
// B rows are transformed in place, twice, into the left factors of the two rank updates: h2 -> ab2 -> ab1,
// U2 -> Ub2 -> Ub1.
// ---------------------------------------------------------------------------------------------------------
template <class K>
GPF_DEV void sf2_reverse_eval(const float* __restrict__ W, const float* __restrict__ Abuf, float* __restrict__ Bbuf,
                              int lane, float ts, float bnd, float vb0, float vb1, float& xbar0, float& xbar1,
                              float (&acc)[16], float (&accw)[16], float (&w3b)[16], float& b3b) {
  constexpr int H1 = K::H1, H2 = K::H2;
  const float* Arow = Abuf + lane * K::AS;
  float* Brow = Bbuf + lane * K::BS;
  const float4 a0 = lds4(Arow);
  const float x0 = a0.x, x1 = a0.y, psi = a0.z, g0 = a0.w, g1 = Brow[15];
  // ---- A.3 (reverse of the closed form)
  const float s0 = fmaf(x0, x0, -bnd), s1 = fmaf(x1, x1, -bnd);
  const float q0 = fmaf(2.0f * psi, x1, s1 * g1);
  const float q1 = -fmaf(2.0f * psi, x0, s0 * g0);
  const float qb0 = vb0 * s0, qb1 = vb1 * s1;
  const float psib = 2.0f * (x1 * qb0 - x0 * qb1);
  const float gb1 = s1 * qb0, gb0 = -s0 * qb1;
  const float sb0 = fmaf(vb0, q0, -g0 * qb1), sb1 = fmaf(vb1, q1, g1 * qb0);
  float xb0 = fmaf(2.0f * x0, sb0, -2.0f * psi * qb1);
  float xb1 = fmaf(2.0f * x1, sb1, 2.0f * psi * qb0);
  b3b += psib;
  // ---- output row + activation 2 (unit k unrolled), in place: B row <- [ab2 | Ub2_0 | Ub2_1]
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const float4 hc = lds4(Brow + 4 * c), p0 = lds4(Brow + 16 + 4 * c), p1 = lds4(Brow + 32 + 4 * c);
    const float4 w3v = ldw4(W, K::W3C + 4 * c);
    const float hh[4] = {hc.x, hc.y, hc.z, hc.w}, uu0[4] = {p0.x, p0.y, p0.z, p0.w}, uu1[4] = {p1.x, p1.y, p1.z, p1.w};
    const float w3[4] = {w3v.x, w3v.y, w3v.z, w3v.w};
    float ab[4], ub0[4], ub1[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int k = 4 * c + i;
      if (k < H2) {
        const float h = hh[i], sp = fmaf(-h, h, 1.0f);
        const float tb = fmaf(gb0, uu0[i], gb1 * uu1[i]);             // sum_j Tb2[k][j] U2[k][j] / w3[k]
        w3b[k] = fmaf(psib, h, fmaf(sp, tb, w3b[k]));                 // psib h2 + sum_j gb_j T2[k][j]
        const float ws = w3[i] * sp;
        ab[i] = ws * fmaf(-2.0f * h, tb, psib);                       // hb2 sp2 + rs sigma''
        ub0[i] = ws * gb0;
        ub1[i] = ws * gb1;
      } else {
        ab[i] = ub0[i] = ub1[i] = 0.0f;
      }
    }
    sts4(Brow + 4 * c, ab[0], ab[1], ab[2], ab[3]);
    sts4(Brow + 16 + 4 * c, ub0[0], ub0[1], ub0[2], ub0[3]);
    sts4(Brow + 32 + 4 * c, ub1[0], ub1[1], ub1[2], ub1[3]);
  }
  __syncwarp();
  // ---- dW2[k][m] += sum_s ab2_s[k] h1_s[m] + sum_j Ub2_s[k][j] T1_s[m][j]: lane (slice, ti, tj) owns the 4x4 tile
  //      (4ti.., 4tj..) over the samples s = slice (mod 2); T1 = (1 - h1^2) W1x is rebuilt from the h1 chunk
  {
    const int tj = lane & 3, ti = (lane >> 2) & 3, slice = lane >> 4;
    float wa[4], wb[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 w1 = ldsw4(W + K::W1P + 4 * (4 * tj + i));  // lane-dependent: shared-memory copy
      wa[i] = w1.x;
      wb[i] = w1.y;
    }
    const float* ap = Abuf + slice * K::AS + 4 + 4 * tj;
    const float* bp = Bbuf + slice * K::BS + 4 * ti;
    float2 wap[2] = {make_float2(wa[0], wa[1]), make_float2(wa[2], wa[3])};
    float2 wbp[2] = {make_float2(wb[0], wb[1]), make_float2(wb[2], wb[3])};
    float2 accp[4][2];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      accp[r][0] = make_float2(acc[4 * r], acc[4 * r + 1]);
      accp[r][1] = make_float2(acc[4 * r + 2], acc[4 * r + 3]);
    }
    const float2 one2 = make_float2(1.0f, 1.0f);
#pragma unroll 2
    for (int q = 0; q < 16; ++q) {
      const float4 hc = lds4(ap + q * (2 * K::AS));
      const float4 l0 = lds4(bp + q * (2 * K::BS)), l1 = lds4(bp + q * (2 * K::BS) + 16), l2 = lds4(bp + q * (2 * K::BS) + 32);
      float2 r0[2] = {make_float2(hc.x, hc.y), make_float2(hc.z, hc.w)}, r1[2], r2[2];
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const float2 sp = __ffma2_rn(make_float2(-r0[i].x, -r0[i].y), r0[i], one2);
        r1[i] = __fmul2_rn(sp, wap[i]);
        r2[i] = __fmul2_rn(sp, wbp[i]);
      }
      const float la[4] = {l0.x, l0.y, l0.z, l0.w}, lb[4] = {l1.x, l1.y, l1.z, l1.w}, lc[4] = {l2.x, l2.y, l2.z, l2.w};
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const float2 lad = make_float2(la[r], la[r]), lbd = make_float2(lb[r], lb[r]), lcd = make_float2(lc[r], lc[r]);
#pragma unroll
        for (int i = 0; i < 2; ++i)
          accp[r][i] = __ffma2_rn(lad, r0[i], __ffma2_rn(lbd, r1[i], __ffma2_rn(lcd, r2[i], accp[r][i])));
      }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      acc[4 * r] = accp[r][0].x; acc[4 * r + 1] = accp[r][0].y;
      acc[4 * r + 2] = accp[r][1].x; acc[4 * r + 3] = accp[r][1].y;
    }
  }
  // ---- back through W2 (rolled over the output unit k): hb1 = W2^T ab2, Tb1[:,j] = W2^T Ub2[:,j]
  float2 hbp[8], t0p[8], t1p[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) hbp[j] = t0p[j] = t1p[j] = make_float2(0.0f, 0.0f);
  GPF_SF2_LOOP
  for (int c = 0; c < 4; ++c) {
    const float4 abv = lds4(Brow + 4 * c), p0 = lds4(Brow + 16 + 4 * c), p1 = lds4(Brow + 32 + 4 * c);
    const float ab[4] = {abv.x, abv.y, abv.z, abv.w}, ub0[4] = {p0.x, p0.y, p0.z, p0.w}, ub1[4] = {p1.x, p1.y, p1.z, p1.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int k = 4 * c + i;
      const float2 abd = make_float2(ab[i], ab[i]), d0 = make_float2(ub0[i], ub0[i]), d1 = make_float2(ub1[i], ub1[i]);
      float2 wp[8];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 t = ldw4(W, K::W2R + 16 * k + 4 * q);
        wp[2 * q] = make_float2(t.x, t.y);
        wp[2 * q + 1] = make_float2(t.z, t.w);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        hbp[j] = __ffma2_rn(wp[j], abd, hbp[j]);
        t0p[j] = __ffma2_rn(wp[j], d0, t0p[j]);
        t1p[j] = __ffma2_rn(wp[j], d1, t1p[j]);
      }
    }
  }
  float hb[16], tb0[16], tb1[16];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    hb[2 * j] = hbp[j].x; hb[2 * j + 1] = hbp[j].y;
    tb0[2 * j] = t0p[j].x; tb0[2 * j + 1] = t0p[j].y;
    tb1[2 * j] = t1p[j].x; tb1[2 * j + 1] = t1p[j].y;
  }
  // ---- activation 1 + layer 1 (unit m unrolled); B row <- [ab1 | Ub1_0 | Ub1_1]
  __syncwarp();  // every lane is done reading the other lanes' [ab2 | Ub2] rows
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const float4 hc = lds4(Arow + 4 + 4 * c);
    const float hh[4] = {hc.x, hc.y, hc.z, hc.w};
    float ab[4], ub0[4], ub1[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int m = 4 * c + i;
      if (m < H1) {
        const float4 w1 = ldw4(W, K::W1P + 4 * m);
        const float h = hh[i], sp = fmaf(-h, h, 1.0f);
        const float rs = fmaf(tb0[m], w1.x, tb1[m] * w1.y);
        ab[i] = sp * fmaf(-2.0f * h, rs, hb[m]);
        ub0[i] = sp * tb0[m];
        ub1[i] = sp * tb1[m];
        xb0 = fmaf(w1.x, ab[i], xb0);
        xb1 = fmaf(w1.y, ab[i], xb1);
      } else {
        ab[i] = ub0[i] = ub1[i] = 0.0f;
      }
    }
    sts4(Brow + 4 * c, ab[0], ab[1], ab[2], ab[3]);
    sts4(Brow + 16 + 4 * c, ub0[0], ub0[1], ub0[2], ub0[3]);
    sts4(Brow + 32 + 4 * c, ub1[0], ub1[1], ub1[2], ub1[3]);
  }
  __syncwarp();
  // ---- dW1[m][:] += sum_s ab1_s[m] [x0, x1, t, 1]_s + [Ub1_s[m][0], Ub1_s[m][1], 0, 0]: lane (slice, ti) owns rows
  //      4ti.. over the samples s = slice (mod 8)
  {
    const int ti = lane & 3, slice = lane >> 2;
    const float* ap = Abuf + slice * K::AS;
    const float* bp = Bbuf + slice * K::BS + 4 * ti;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 xa = lds4(ap + q * (8 * K::AS));
      const float4 l0 = lds4(bp + q * (8 * K::BS)), l1 = lds4(bp + q * (8 * K::BS) + 16), l2 = lds4(bp + q * (8 * K::BS) + 32);
      const float la[4] = {l0.x, l0.y, l0.z, l0.w}, lb[4] = {l1.x, l1.y, l1.z, l1.w}, lc[4] = {l2.x, l2.y, l2.z, l2.w};
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        accw[4 * r + 0] += fmaf(la[r], xa.x, lb[r]);
        accw[4 * r + 1] += fmaf(la[r], xa.y, lc[r]);
        accw[4 * r + 2] = fmaf(la[r], ts, accw[4 * r + 2]);
        accw[4 * r + 3] += la[r];
      }
    }
  }
  xbar0 = xb0;
  xbar1 = xb1;
}

}  // namespace gpflow
