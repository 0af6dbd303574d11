// gpflow_b200 -- sf2e kernel proper + host launch: the sf2r train step / backward op WITH the Euler penalization
// (res_net.py:291-302,319-358).  Per RK4 step: 4 plain evaluations (sf2r), then at (x_{k+1}, t_k) 4 dual evaluations
// (directions w and z, points x and x + delta v) whose epsilon parts give q = z . J_a w - w . J_a z; the reverse sweep
// walks the same 8 stash rows per step backwards.  See gpflow_sf2e.cuh.
#pragma once
#include "gpflow_sf2r_kernel.cuh"
#include "gpflow_sf2e_rev.cuh"

namespace gpflow {

struct SF2EArgs {
  SF2Args b;
  const float* zw;        // [N][2][2]: z then w of every step (res_net.py:330-333)
  const float* reg_cot;   // nullable: cotangent of the regularization value (backward op)
  float* qbuf;            // [nslots][N][32]: q of every sample of the tile in flight, forward -> reverse sweep
  float euler_coef;
};

// 8 warps: 255 registers per thread (no spills), 28 KB of shared memory per warp (two 8 KB stash regions + work rows).
// Measured alternative (profiles/r2_sf2e_ncu_summary.txt): ONE stash region (plain evaluations alternate between its halves,
// dual rows single-buffered) and 84-float work rows fit 12 warps at 168 registers: 155.9 vs 157.6 ms on the C4 Euler row and
// slower at small batches -- 750 B of spills, and with 12 warps spread over 7 000 instructions of code `no_instruction` is
// back at 0.62 stall cycles per issue (0.18 with 8 warps).  Not kept.
// warps per CTA (= per SM).  8: two stash regions per warp, 255 registers.  9 - 11 (GPF_SF2E_WARPS): ONE region per warp, and the
// register file then allows 168 registers (472 B of spills).  Measured with 11 warps on the phase-aligned build, C4 Euler
// row: 114.5 ms against 91.7 ms with 8 warps (parity green) -- the dual evaluations need their 255 registers more than the
// schedulers need a third warp.
#ifndef GPF_SF2E_WARPS
#define GPF_SF2E_WARPS 8
#endif
constexpr int kSF2EWarps = GPF_SF2E_WARPS;

template <int H1, int H2, int NW>
__global__ void __launch_bounds__(NW * 32, 1) gpflow_sf2e_kernel(const SF2EArgs ea) {
  typedef SF2E<H1, H2, NW> K;
  typedef typename K::R R;
  const SF2Args& a = ea.b;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint64_t bar_stage;
  __shared__ uint64_t bars[NW][2];
  float* const W = reinterpret_cast<float*>(smem_raw);
  float* const sct = W + R::WEND;
  float* const times = sct + R::SCT;
  float* const raw = times + round_up(a.n_steps + 1, 4);
  float* const warp_area = raw + round_up(R::NPARAM, 4);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, l7 = lane & 7;

  stage_params_to_smem<float>(raw, a.params, R::NPARAM, &bar_stage);
  for (int i = tid; i < R::WEND; i += NW * 32) W[i] = sf2r_pack_value<R>(raw, i);
  if (tid < 4) {
    const float hdt_ = a.dt / 2.0f, c6_ = a.dt / 6.0f;
    const float wk_ = (tid == 1 || tid == 2) ? 2.0f : 1.0f;
    sts4(sct + 4 * tid, tid == 0 ? 0.0f : (tid == 3 ? a.dt : hdt_), wk_, tid == 3 ? 0.0f : (tid == 2 ? a.dt : hdt_),
         wk_ * c6_);
  }
  if (tid == 0) {
    float t = a.t0;
    for (int i = 0; i <= a.n_steps; ++i) {
      times[i] = t;
      t = t + a.dt;
    }
  }
  if (lane == 0) {
    mbar_init(&bars[warp][0], 1);
    mbar_init(&bars[warp][1], 1);
    fence_barrier_init();
  }
  float* const wa = warp_area + (size_t)warp * K::WARP_FL;
  float* const Reg0 = wa;                  // stash region r = Reg0 + r * REG_FL (plain rows use its first 4 KB)
  float* const Wk = wa + K::NREG * K::REG_FL;    // work rows (plain: [32][68], dual: [32][92])
  const int slot = warp * gridDim.x + blockIdx.x, nslots = gridDim.x * NW;
  constexpr int G_PEN = R::NACC, G_W2D = R::NACC + 1, NACC = R::NACC + 17;   // + penalty, + the dual evaluations' dW2 tile
  float* const G = a.g2 + (size_t)slot * NACC * 32;
  for (int i = 0; i < NACC; ++i) G[i * 32 + lane] = 0.0f;
  __syncthreads();

  const bool wrap = (a.flags & 0x1u) != 0;
  const float bnd = 1.0f - Eps<float>::value;
  const float c6 = a.dt / 6.0f;
  const float delta = (float)(2.0 * sqrt((double)Eps<float>::value));  // res_net.py:346
  const float rdelta = 1.0f / delta;
  const float rsq2s = 1.0f / (1.4142135623730951f * a.sigma), sq2s = 1.4142135623730951f * a.sigma;
  const int N = a.n_steps;
  char* const stash = a.stash + (size_t)slot * N * K::STEP_BYTES;
  float* const qrow = ea.qbuf + (size_t)slot * N * 32 + lane;
  const float pen_w = ea.euler_coef * a.loss_scale;
  float reg_cot = 1.0f;
  if (a.x_data == nullptr && ea.reg_cot != nullptr) reg_cot = *ea.reg_cot;   // backward op
  uint32_t phbits = 0;
  uint64_t* const mybars = &bars[warp][0];

  float acc[16], accd[16], accw[16], acc3[4], b3b = 0.0f;   // accd: dW2 tile of the dual evaluations (their own lane layout)
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = accd[i] = accw[i] = 0.0f;
#pragma unroll
  for (int i = 0; i < 4; ++i) acc3[i] = 0.0f;

  const long long ntiles = (a.B + 31) / 32;
  if (a.stagger_cycles != 0) {   // warp w starts (w * 5 mod NW) / NW * |stagger_cycles| late, see gpflow_sf2r_kernel
    const long long wait = (long long)(a.stagger_cycles < 0 ? -a.stagger_cycles : a.stagger_cycles) * ((warp * 5) % NW) / NW;
    const long long t0 = clock64();
    while (clock64() - t0 < wait) __nanosleep(256);
  }
  for (long long tile = slot; tile < ntiles; tile += nslots) {
    const long long b = tile * 32 + lane;
    const bool valid = b < a.B;
    float xin0 = 0.f, xin1 = 0.f;
    if (valid) {
      xin0 = a.x_in[2 * b];
      xin1 = a.x_in[2 * b + 1];
    }
    float x0 = wrap ? erf_real(xin0 * rsq2s) : xin0, x1 = wrap ? erf_real(xin1 * rsq2s) : xin1;
    // The CTA's warps run the same 2 N steps per tile on different samples.  A named barrier over the warps that have a tile
    // in this round (a prefix of the warps: slot = warp * grid + block) keeps them within one RK4 step of each other, so that
    // they share the instruction-cache footprint of the 7 000-instruction kernel (the generic kernel does the same).
    long long n_act = (ntiles - (tile - slot) - blockIdx.x + gridDim.x - 1) / gridDim.x;
    if (n_act > NW) n_act = NW;
    const unsigned n_act_threads = 32u * (unsigned)n_act;
    auto cta_align = [&]() {
      if (a.cta_align && n_act_threads > 32u) asm volatile("bar.sync 1, %0;" ::"r"(n_act_threads) : "memory");
    };

    // ---------------- forward sweep ----------------
    float pen = 0.0f;
    int op = 0;   // stash operations issued for this tile (region = op & 1)
    auto begin_op = [&]() -> float* {
      if (K::SINGLE) {
        if (op >= 1) {  // the bulk store of the previous operation must have read the region
          if (lane == 0) bulk_wait_read<0>();
          __syncwarp();
        }
        return Reg0;
      }
      if (op >= 2) {  // the bulk store that read this region two operations ago must have drained
        if (lane == 0) bulk_wait_read<1>();
        __syncwarp();
      }
      return Reg0 + (op & 1) * K::REG_FL;
    };
    auto end_op = [&](const float* reg, int step, int j) {
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) {
        bulk_s2g(stash + K::op_offset(step, j), reg, K::op_bytes(j));
        bulk_commit();
      }
      ++op;
    };
#pragma unroll 1
    for (int i = 0; i < N; ++i) {
      cta_align();
      const float t = times[i];
      float ys0 = x0, ys1 = x1, ks0 = 0.f, ks1 = 0.f;
#pragma unroll 1
      for (int sg = 0; sg < 4; ++sg) {
        const float4 sc = ldsw4(sct + 4 * sg);
        float* const reg = begin_op();
        float v0, v1;
        sf2r_forward_eval<R>(W, reg + lane * R::SROW, l7, ys0, ys1, t + sc.x, bnd, v0, v1);
        end_op(reg, i, sg);
        ks0 = fmaf(sc.y, v0, ks0);
        ks1 = fmaf(sc.y, v1, ks1);
        ys0 = fmaf(sc.z, v0, x0);
        ys1 = fmaf(sc.z, v1, x1);
      }
      x0 = fmaf(c6, ks0, x0);
      x1 = fmaf(c6, ks1, x1);
      // ---- Euler penalization at (x_{i+1}, t_i): q = z . J_a w - w . J_a z   (res_net.py:319-358)
      const float4 zw = make_float4(ea.zw[4 * i], ea.zw[4 * i + 1], ea.zw[4 * i + 2], ea.zw[4 * i + 3]);   // z0 z1 w0 w1
      float q = 0.0f;
      du P0 = d_make(0.0f, 0.0f), P1 = P0, V10 = P0, V11 = P0, V0 = P0, V1 = P0;
#pragma unroll 1
      for (int e = 0; e < 4; ++e) {   // ONE call site of the dual evaluation: (direction w | z) x (point x | x + delta v)
        const bool second = (e & 1) != 0, dirz = e >= 2;
        if (!second) {
          P0 = d_make(x0, dirz ? zw.x : zw.z);
          P1 = d_make(x1, dirz ? zw.y : zw.w);
        } else {
          V10 = V0; V11 = V1;
          P0 = d_fma(delta, V0, P0);
          P1 = d_fma(delta, V1, P1);
        }
        float* const reg = begin_op();
        sf2e_forward_eval<K>(W, reg + lane * K::DROW, Wk + lane * K::WS, l7, P0, P1, second ? t + delta : t, bnd, V0, V1);
        end_op(reg, i, 4 + e);
        if (second) {
          const float a0 = (V0.y - V10.y) * rdelta, a1 = (V1.y - V11.y) * rdelta;   // eps part of a along the direction
          q += dirz ? -fmaf(zw.z, a0, zw.w * a1) : fmaf(zw.x, a0, zw.y * a1);      // q = z . a_w - w . a_z
        }
      }
      if (!valid) q = 0.0f;
      pen = fmaf(q, q, pen);
      qrow[i * 32] = q;
    }
    // ---------------- wrapper epilogue (network.py:178-185) ----------------
    float y0 = x0, y1 = x1;
    bool in0 = true, in1 = true;
    if (wrap) {
      const float lim = 1.0f - Eps<float>::value;
      in0 = (x0 >= -lim) && (x0 <= lim);
      in1 = (x1 >= -lim) && (x1 <= lim);
      y0 = sq2s * erfinv_real(fminf(fmaxf(x0, -lim), lim));
      y1 = sq2s * erfinv_real(fminf(fmaxf(x1, -lim), lim));
    }
    if (valid && a.y_out != nullptr) {
      a.y_out[2 * b] = y0;
      a.y_out[2 * b + 1] = y1;
    }
    // ---------------- seed of the reverse sweep ----------------
    float yb0 = 0.f, yb1 = 0.f, ot = 0.f;
    if (valid) {
      if (a.x_data != nullptr) {
        const float d0 = a.x_data[2 * b] - y0, d1 = a.x_data[2 * b + 1] - y1;
        ot = fmaf(d0, d0, d1 * d1);
        yb0 = -2.0f * d0 * a.loss_scale;
        yb1 = -2.0f * d1 * a.loss_scale;
      } else {
        yb0 = a.grad_y[2 * b];
        yb1 = a.grad_y[2 * b + 1];
      }
    }
    float xb0 = yb0, xb1 = yb1;
    if (wrap) {
      const float i2s = 1.0f / (2.0f * a.sigma * a.sigma);
      xb0 = in0 ? yb0 * (1.2533141373155001f * a.sigma * exp_real(y0 * y0 * i2s)) : 0.0f;
      xb1 = in1 ? yb1 * (1.2533141373155001f * a.sigma * exp_real(y1 * y1 * i2s)) : 0.0f;
    }
    // ---------------- reverse sweep: stash operations 8N-1 ... 0 ----------------
    auto issue_load = [&](int o) {   // lane 0 only
      const int r = K::SINGLE ? 0 : (o & 1), j = o & 7;
      mbar_expect_tx(mybars + r, K::op_bytes(j));
      bulk_g2s(Reg0 + r * K::REG_FL, stash + K::op_offset(o >> 3, j), K::op_bytes(j), mybars + r);
    };
    auto next_rows = [&](int o) -> float* {   // rows of operation o: prefetch o - 1, wait for o
      fence_proxy_async();
      __syncwarp();
      if (K::SINGLE) {   // the previous operation's rows are dead now: load this one's into the same region
        if (lane == 0 && o != 8 * N - 1) issue_load(o);
        mbar_wait(mybars, phbits & 1u);
        phbits ^= 1u;
        return Reg0;
      }
      if (o > 0 && lane == 0) issue_load(o - 1);
      const int r = o & 1;
      mbar_wait(mybars + r, (phbits >> r) & 1u);
      phbits ^= 1u << r;
      return Reg0 + r * K::REG_FL;
    };
    if (lane == 0) {
      bulk_wait_all<0>();
      fence_proxy_async_all();
      issue_load(8 * N - 1);
    }
#pragma unroll 1
    for (int i = N - 1; i >= 0; --i) {
      cta_align();
      const float t = times[i];
      // ---- penalty: d(pen_w q^2) at (x_{i+1}, t_i); operations 7, 6 (direction z) then 5, 4 (direction w)
      {
        const float4 zw = make_float4(ea.zw[4 * i], ea.zw[4 * i + 1], ea.zw[4 * i + 2], ea.zw[4 * i + 3]);
        const float g = 2.0f * qrow[i * 32] * pen_w * reg_cot;
        du Yb0 = d_make(0.0f, 0.0f), Yb1 = Yb0;
        float u0 = 0.0f, u1 = 0.0f;
#pragma unroll 1
        for (int e = 3; e >= 0; --e) {   // ONE call site of the dual reverse sweep
          const bool second = (e & 1) != 0, dirz = e >= 2;
          du S0, S1;
          if (second) {   // direction w: cotangent of a_w is g z;  direction z: cotangent of a_z is -g w
            u0 = (dirz ? -g * zw.z : g * zw.x) * rdelta;
            u1 = (dirz ? -g * zw.w : g * zw.y) * rdelta;
            S0 = d_make(u0, 0.0f);
            S1 = d_make(u1, 0.0f);
          } else {
            S0 = d_fma(delta, Yb0, d_make(-u0, 0.0f));
            S1 = d_fma(delta, Yb1, d_make(-u1, 0.0f));
          }
          float* const rows = next_rows(8 * i + 4 + e);
          du R0, R1;
          sf2e_reverse_eval<K>(W, rows, Wk, lane, second ? t + delta : t, bnd, S0, S1, R0, R1, accd, accw, acc3, b3b);
          xb0 += R0.y;
          xb1 += R1.y;
          Yb0 = R0;
          Yb1 = R1;
        }
      }
      // ---- the RK4 step (A.5)
      float sum0 = 0.f, sum1 = 0.f, prev0 = 0.f, prev1 = 0.f;
#pragma unroll 1
      for (int sg = 3; sg >= 0; --sg) {
        const float4 sc = ldsw4(sct + 4 * sg);
        const float* rows = next_rows(8 * i + sg);
        const float kb0 = fmaf(sc.z, prev0, sc.w * xb0), kb1 = fmaf(sc.z, prev1, sc.w * xb1);
        float r0, r1;
        sf2r_reverse_eval<R>(W, rows, Wk, lane, t + sc.x, bnd, kb0, kb1, r0, r1, acc, accw, acc3, b3b);
        prev0 = r0; prev1 = r1;
        sum0 += r0; sum1 += r1;
      }
      xb0 += sum0;
      xb1 += sum1;
    }
    if (valid && a.grad_x_out != nullptr) {
      float gx0 = xb0, gx1 = xb1;
      if (wrap) {
        const float i2s = 1.0f / (2.0f * a.sigma * a.sigma), c = 0.7978845608028654f / a.sigma;
        gx0 *= c * exp_real(-xin0 * xin0 * i2s);
        gx1 *= c * exp_real(-xin1 * xin1 * i2s);
      }
      a.grad_x_out[2 * b] = gx0;
      a.grad_x_out[2 * b + 1] = gx1;
    }
    // ---------------- per-tile flush into the second-level accumulators ----------------
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      G[(R::G_W2 + i) * 32 + lane] += acc[i];
      G[(G_W2D + i) * 32 + lane] += accd[i];
      G[(R::G_W1 + i) * 32 + lane] += accw[i];
      acc[i] = accd[i] = accw[i] = 0.0f;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      G[(R::G_W3 + i) * 32 + lane] += acc3[i];
      acc3[i] = 0.0f;
    }
    G[R::G_B3 * 32 + lane] += b3b;
    b3b = 0.0f;
    G[R::G_OT * 32 + lane] += ot;
    G[G_PEN * 32 + lane] += pen;
  }

  // ---------------- per-warp partial vector [ot, reg, lagr | grads] ----------------
  __syncwarp();
  float* part = a.partials + (size_t)slot * a.npart;
  const float ot_w = warp_sum(G[R::G_OT * 32 + lane]);
  const float pen_ws = warp_sum(G[G_PEN * 32 + lane]);
  if (lane == 0) {
    part[0] = ot_w * a.loss_scale;
    part[1] = pen_ws * pen_w;
    part[2] = 0.0f;
  }
  for (int e = lane; e < R::NPARAM; e += 32) part[3 + e] = sf2r_partial_entry<R>(G, e, G + G_W2D * 32);
}

// ---- host side --------------------------------------------------------------------------------------------------
template <int H1, int H2>
size_t sf2e_workspace(const GpFlowDesc* d, int64_t B) {
  typedef SF2E<H1, H2, kSF2EWarps> K;
  const int64_t tiles = (B + 31) / 32;
  const int64_t ctas = tiles < kMaxSMs ? (tiles < 1 ? 1 : tiles) : kMaxSMs;
  const int64_t slots = ctas * kSF2EWarps;
  const int64_t stash_slots = tiles < slots ? tiles : slots;
  size_t bytes = 704 * 4;
  bytes += align_up((size_t)slots * (3 + K::R::NPARAM) * 4, 256);
  bytes += align_up((size_t)slots * (K::R::NACC + 17) * 32 * 4, 256);
  bytes += align_up((size_t)slots * d->n_steps * 32 * 4, 256);
  bytes += (size_t)stash_slots * d->n_steps * K::STEP_BYTES;
  return bytes;
}

template <int H1, int H2>
int launch_sf2e(const GpFlowDesc* d, const Ptrs<float>& p, int64_t B, void* ws, size_t ws_bytes, cudaStream_t st,
                int sm_count) {
  constexpr int NW = kSF2EWarps;
  typedef SF2E<H1, H2, NW> K;
  typedef typename K::R R;
  auto kern = gpflow_sf2e_kernel<H1, H2, NW>;
  const size_t smem = K::smem_bytes(d->n_steps);
  if (smem > 227 * 1024) return GPFLOW_ERR_UNSUPPORTED;
  if (p.zw == nullptr) return GPFLOW_ERR_NULL_POINTER;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
    cudaGetLastError();
    return GPFLOW_ERR_CUDA;
  }
  if (ws == nullptr || ws_bytes < sf2e_workspace<H1, H2>(d, B)) return GPFLOW_ERR_WORKSPACE;
  if (reinterpret_cast<uintptr_t>(ws) & 255) return GPFLOW_ERR_MISALIGNED;
  const int64_t tiles = (B + 31) / 32;
  int64_t grid = tiles;
  if (grid > sm_count) grid = sm_count;
  if (grid < 1) grid = 1;
  const int64_t slots = grid * NW;
  const int npart = 3 + R::NPARAM;

  SF2EArgs ea;
  memset(&ea, 0, sizeof(ea));
  SF2Args& a = ea.b;
  a.params = p.params; a.x_in = p.x_in; a.x_data = p.x_data; a.grad_y = p.grad_y;
  a.y_out = p.y_out; a.grad_x_out = p.grad_x_out;
  char* wsp = reinterpret_cast<char*>(ws);
  float* packed = reinterpret_cast<float*>(wsp);
  wsp += 704 * 4;
  a.partials = reinterpret_cast<float*>(wsp);
  wsp += align_up((size_t)slots * npart * 4, 256);
  a.g2 = reinterpret_cast<float*>(wsp);
  wsp += align_up((size_t)slots * (R::NACC + 17) * 32 * 4, 256);
  ea.qbuf = reinterpret_cast<float*>(wsp);
  wsp += align_up((size_t)slots * d->n_steps * 32 * 4, 256);
  a.stash = wsp;
  a.B = B; a.n_steps = d->n_steps; a.npart = npart; a.flags = d->flags;
  const bool rev = (d->flags & GPFLOW_FLAG_REVERSE) != 0;
  a.dt = rev ? -(float)d->dt : (float)d->dt;
  a.t0 = rev ? (float)(d->dt * (double)d->n_steps_nominal) : 0.0f;
  a.sigma = (float)d->sigma;
  a.loss_scale = (float)d->loss_scale;
  ea.zw = p.zw;
  ea.reg_cot = p.reg_cot;
  ea.euler_coef = (float)d->euler_coef;
  // staggered warp start (GPFLOW_SF2E_STAGGER = cycles spread over the 8 warps; default 0 = off).  Measured on the C4 Euler row
  // (tools/sf2e_stagger_ab.sh, profiles/r2_sf2e_stagger_and_batch_sweep.txt): 0 / 4e4 / 1.6e5 / 6.4e5 / 2.5e6 cycles ->
  // 141.5 / 141.0 / 147.0 / 144.6 / 138.5 ms, i.e. run-to-run noise: unlike sf2r the 8 warps do not gain from being out of phase.
  static const int stagger_env = [] { const char* e = getenv("GPFLOW_SF2E_STAGGER"); return e ? atoi(e) : 0; }();
  a.stagger_cycles = tiles >= 4 * slots ? stagger_env : 0;
  // phase alignment of the CTA's warps (GPFLOW_SF2E_ALIGN=0: free-running).  Measured on the C4 Euler row, same box, alternating:
  // 138.7 / 139.9 ms free-running, 103.2 / 103.2 ms aligned (2.99e7 -> 4.06e7 samples/s).  The row also ran 16 % slower per
  // sample at 2^22 than at 2^19 samples (profiles/r2_sf2e_stagger_and_batch_sweep.txt): free-running warps drift apart over
  // a long launch and each then misses the instruction cache on its own (7 080 SASS instructions = 113 KB).
  static const int align_env = [] { const char* e = getenv("GPFLOW_SF2E_ALIGN"); return e ? atoi(e) : 1; }();
  a.cta_align = align_env;
  const int rc = sf2r_launch_guarded<R>(p.params, packed, st, [&]() { kern<<<(unsigned)grid, NW * 32, smem, st>>>(ea); });
  if (rc != GPFLOW_OK) return rc;
  gpflow_finalize_wide_kernel<float><<<(npart + 31) / 32, 256, 0, st>>>(a.partials, (int)slots, npart, npart, p.scalars_out,
                                                                        p.grads_out);
  if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  return GPFLOW_OK;
}

}  // namespace gpflow
