// gpflow_b200 -- sf2 kernel proper + host launch.  See gpflow_sf2.cuh for the design.
#pragma once
#include <stdlib.h>
#include <string.h>

#include "gpflow_dispatch.h"
#include "gpflow_sf2_eval.cuh"

namespace gpflow {

// packed weight i (layout in SF2<>) from the flat nn.Linear parameters: transpose / collapse / pad
template <class K>
GPF_DEV float sf2_pack_value(const float* __restrict__ raw, int i) {
  constexpr int H1 = K::H1, H2 = K::H2;
  float v = 0.0f;
  if (i < K::W2T) {
    const int k = i >> 2, c = i & 3;
    if (k < H1) v = c < 3 ? raw[K::OFF_W1 + 3 * k + c] : raw[K::OFF_B1 + k];
  } else if (i < K::W2R) {
    const int m = (i - K::W2T) >> 4, k = (i - K::W2T) & 15;
    if (k < H2) v = m < H1 ? raw[K::OFF_W2 + k * H1 + m] : (m == 15 ? raw[K::OFF_B2 + k] : 0.0f);
  } else if (i < K::W3C) {
    const int k = (i - K::W2R) >> 4, m = (i - K::W2R) & 15;
    if (k < H2 && m < H1) v = raw[K::OFF_W2 + k * H1 + m];
  } else if (i < K::B3C) {
    const int k = i - K::W3C;
    if (k < H2) v = raw[K::OFF_W3 + k] - raw[K::OFF_W3 + H2 + k];
  } else if (i == K::B3C) {
    v = raw[K::OFF_B3] - raw[K::OFF_B3 + 1];
  }
  return v;
}

template <class K>
__global__ void sf2_pack_kernel(const float* __restrict__ params, float* __restrict__ packed) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < 640) packed[i] = i < K::WEND ? sf2_pack_value<K>(params, i) : 0.0f;
}

template <int H1, int H2, int NW, bool TRAIN>
__global__ void __launch_bounds__(NW * 32, 1) gpflow_sf2_kernel(const SF2Args a) {
  typedef SF2<H1, H2, NW> K;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint64_t bar_stage;
  __shared__ uint64_t bars[NW][2];
  float* const W = reinterpret_cast<float*>(smem_raw);
  float* const sct = W + K::WEND;  // [4][4] stage constants
  float* const times = sct + K::SCT;
  float* const raw = times + round_up(a.n_steps + 1, 4);
  float* const warp_area = raw + round_up(K::NPARAM, 4);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // ---- prologue: TMA bulk copy of the flat parameters, re-pack (transpose / collapse / pad), time table ----
  stage_params_to_smem<float>(raw, a.params, K::NPARAM, &bar_stage);
  for (int i = tid; i < K::WEND; i += NW * 32) W[i] = sf2_pack_value<K>(raw, i);
  if (tid < 4) {
    // RK4 stage sg (ode_order_method.py:66-71): time offset, weight of k_sg in the sum, step to the next stage point
    // (0 after the last stage; this is also the adjoint's dt-factor of yb_{sg+1}), weight * dt/6
    const float hdt_ = a.dt / 2.0f, c6_ = a.dt / 6.0f;
    const float wk_ = (tid == 1 || tid == 2) ? 2.0f : 1.0f;
    sts4(sct + 4 * tid, tid == 0 ? 0.0f : (tid == 3 ? a.dt : hdt_), wk_, tid == 3 ? 0.0f : (tid == 2 ? a.dt : hdt_),
         wk_ * c6_);
  }
  if (tid == 0) {
    float t = a.t0;  // repeated addition in the working precision (ode_solver.py:80, res_net.py:262)
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
  float* const Abuf0 = wa;                         // A(b) = Abuf0 + b * A_FL
  float* const Bbuf0 = wa + 2 * K::A_FL;           // B(b) = Bbuf0 + b * B_FL
  // warp-major slot numbering: a small batch (tiles < resident warps) spreads over the SMs first, then over the
  // warps of an SM
  const int slot = warp * gridDim.x + blockIdx.x, nslots = gridDim.x * NW;
  float* const G = a.g2 + (size_t)slot * K::NACC * 32;
  if (TRAIN) {
    for (int i = 0; i < K::NACC; ++i) G[i * 32 + lane] = 0.0f;
  }
  __syncthreads();

  const bool wrap = (a.flags & 0x1u) != 0;
  const float bnd = 1.0f - Eps<float>::value;  // velocity_field.py:96
  const float c6 = a.dt / 6.0f;
  const float rsq2s = 1.0f / (1.4142135623730951f * a.sigma), sq2s = 1.4142135623730951f * a.sigma;
  const int E = 4 * a.n_steps;
  char* const stash = a.stash + (size_t)slot * E * K::STASH_BYTES;
  uint32_t phbits = 0;  // bit b = mbarrier phase parity of stash buffer b
  uint64_t* const mybars = &bars[warp][0];

  float acc[16], accw[16], w3b[16], b3b = 0.0f;
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = accw[i] = w3b[i] = 0.0f;

  const long long ntiles = (a.B + 31) / 32;
  for (long long tile = slot; tile < ntiles; tile += nslots) {
    const long long b = tile * 32 + lane;
    const bool valid = b < a.B;
    float xin0 = 0.f, xin1 = 0.f;
    if (valid) {
      xin0 = a.x_in[2 * b];
      xin1 = a.x_in[2 * b + 1];
    }
    float x0 = wrap ? erf_real(xin0 * rsq2s) : xin0, x1 = wrap ? erf_real(xin1 * rsq2s) : xin1;

    // ---------------- forward sweep: 4N velocity evaluations, classical RK4 (ode_order_method.py:66-71) ----------
    float ys0 = x0, ys1 = x1, ks0 = 0.f, ks1 = 0.f;
#pragma unroll 1
    for (int e = 0; e < E; ++e) {
      const int sg = e & 3, buf = e & 1;
      const float4 sc = ldsw4(sct + 4 * sg);
      const float ts = times[e >> 2] + sc.x;  // t, t + dt/2, t + dt/2, t + dt
      float* const Ab = Abuf0 + buf * K::A_FL;
      float* const Bb = Bbuf0 + buf * K::B_FL;
      if (TRAIN && e >= 2) {  // the bulk store that read this buffer two evaluations ago must have drained
        if (lane == 0) bulk_wait_read<1>();
        __syncwarp();
      }
      float v0, v1;
      sf2_forward_eval<K, TRAIN>(W, Ab + lane * K::AS, Bb + lane * K::BS, ys0, ys1, ts, bnd, v0, v1);
      if (TRAIN) {
        fence_proxy_async();  // generic-proxy row writes -> visible to the async proxy (TMA)
        __syncwarp();
        if (lane == 0) {
          char* dst = stash + (size_t)e * K::STASH_BYTES;
          bulk_s2g(dst, Ab, K::A_BYTES);
          bulk_s2g(dst + K::A_BYTES, Bb, K::B_BYTES);
          bulk_commit();
        }
      }
      // ksum = k1 + 2 k2 + 2 k3 + k4 accumulated left to right like the reference expression
      ks0 = fmaf(sc.y, v0, ks0);
      ks1 = fmaf(sc.y, v1, ks1);
      const bool last = sg == 3;
      const float xn0 = fmaf(c6, ks0, x0), xn1 = fmaf(c6, ks1, x1);
      ys0 = last ? xn0 : fmaf(sc.z, v0, x0);
      ys1 = last ? xn1 : fmaf(sc.z, v1, x1);
      x0 = last ? xn0 : x0;
      x1 = last ? xn1 : x1;
      ks0 = last ? 0.0f : ks0;
      ks1 = last ? 0.0f : ks1;
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
    if (!TRAIN) continue;

    // ---------------- seed of the reverse sweep ----------------
    float yb0 = 0.f, yb1 = 0.f, ot = 0.f;
    if (valid) {
      if (a.x_data != nullptr) {
        const float d0 = a.x_data[2 * b] - y0, d1 = a.x_data[2 * b + 1] - y1;
        ot = fmaf(d0, d0, d1 * d1);
        yb0 = -2.0f * d0 * a.loss_scale;  // d/dy of loss_scale * sum (x - y)^2  (network.py:82)
        yb1 = -2.0f * d1 * a.loss_scale;
      } else {
        yb0 = a.grad_y[2 * b];
        yb1 = a.grad_y[2 * b + 1];
      }
    }
    float xb0 = yb0, xb1 = yb1;
    if (wrap) {  // dy/dg = sqrt(pi/2) sigma exp(y^2 / (2 sigma^2)) inside the clip band, 0 outside
      const float i2s = 1.0f / (2.0f * a.sigma * a.sigma);
      xb0 = in0 ? yb0 * (1.2533141373155001f * a.sigma * exp_real(y0 * y0 * i2s)) : 0.0f;
      xb1 = in1 ? yb1 * (1.2533141373155001f * a.sigma * exp_real(y1 * y1 * i2s)) : 0.0f;
    }
    // every stash write of this tile must have landed before it is read back
    if (lane == 0) {
      bulk_wait_all<0>();
      fence_proxy_async_all();
      mbar_expect_tx(mybars + ((E - 1) & 1), K::STASH_BYTES);
      const char* src = stash + (size_t)(E - 1) * K::STASH_BYTES;
      bulk_g2s(Abuf0 + ((E - 1) & 1) * K::A_FL, src, K::A_BYTES, mybars + ((E - 1) & 1));
      bulk_g2s(Bbuf0 + ((E - 1) & 1) * K::B_FL, src + K::A_BYTES, K::B_BYTES, mybars + ((E - 1) & 1));
    }
    // ---------------- reverse sweep (A.5) ----------------
    float sum0 = 0.f, sum1 = 0.f, prev0 = 0.f, prev1 = 0.f;
#pragma unroll 1
    for (int e = E - 1; e >= 0; --e) {
      const int sg = e & 3, buf = e & 1;
      const float4 sc = ldsw4(sct + 4 * sg);
      const float ts = times[e >> 2] + sc.x;
      // prefetch the rows of evaluation e-1 into the other buffer (its previous user, e+1, is complete)
      fence_proxy_async();
      __syncwarp();
      if (e > 0 && lane == 0) {
        const int nb = buf ^ 1;
        mbar_expect_tx(mybars + nb, K::STASH_BYTES);
        const char* src = stash + (size_t)(e - 1) * K::STASH_BYTES;
        bulk_g2s(Abuf0 + nb * K::A_FL, src, K::A_BYTES, mybars + nb);
        bulk_g2s(Bbuf0 + nb * K::B_FL, src + K::A_BYTES, K::B_BYTES, mybars + nb);
      }
      mbar_wait(mybars + buf, (phbits >> buf) & 1u);
      phbits ^= 1u << buf;
      // kb_4 = dt/6 xb; kb_3 = dt/3 xb + dt yb_4; kb_2 = dt/3 xb + dt/2 yb_3; kb_1 = dt/6 xb + dt/2 yb_2
      const float kb0 = fmaf(sc.z, prev0, sc.w * xb0), kb1 = fmaf(sc.z, prev1, sc.w * xb1);
      float r0, r1;
      sf2_reverse_eval<K>(W, Abuf0 + buf * K::A_FL, Bbuf0 + buf * K::B_FL, lane, ts, bnd, kb0, kb1, r0, r1, acc, accw,
                          w3b, b3b);
      prev0 = r0; prev1 = r1;
      sum0 += r0; sum1 += r1;
      const bool first = sg == 0;
      xb0 = first ? xb0 + sum0 : xb0;
      xb1 = first ? xb1 + sum1 : xb1;
      sum0 = first ? 0.0f : sum0;
      sum1 = first ? 0.0f : sum1;
      prev0 = first ? 0.0f : prev0;
      prev1 = first ? 0.0f : prev1;
    }
    if (valid && a.grad_x_out != nullptr) {
      float gx0 = xb0, gx1 = xb1;
      if (wrap) {  // dg0/dx = sqrt(2/pi)/sigma exp(-x^2/(2 sigma^2))
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
      G[(K::G_W2 + i) * 32 + lane] += acc[i];
      G[(K::G_W3 + i) * 32 + lane] += w3b[i];
      G[(K::G_W1 + i) * 32 + lane] += accw[i];
      acc[i] = w3b[i] = accw[i] = 0.0f;
    }
    G[K::G_B3 * 32 + lane] += b3b;
    b3b = 0.0f;
    G[K::G_OT * 32 + lane] += ot;
  }

  if (!TRAIN) return;
  // ---------------- per-warp partial vector [ot, reg, lagr | grads] ----------------
  __syncwarp();
  float* part = a.partials + (size_t)slot * a.npart;
  const float ot_w = warp_sum(G[K::G_OT * 32 + lane]);
  if (lane == 0) {
    part[0] = ot_w * a.loss_scale;
    part[1] = 0.0f;
    part[2] = 0.0f;
  }
  for (int e = lane; e < K::NPARAM; e += 32) {
    float s = 0.0f;
    if (e < K::OFF_W2) {  // W1 [H1][3] then b1 [H1]  ->  tile entry (m & 3, c) of the lanes (slice, ti = m >> 2)
      const int m = e < K::OFF_B1 ? e / 3 : e - K::OFF_B1, c = e < K::OFF_B1 ? e % 3 : 3;
      for (int sl = 0; sl < 8; ++sl) s += G[(K::G_W1 + (m & 3) * 4 + c) * 32 + sl * 4 + (m >> 2)];
    } else if (e < K::OFF_W3) {  // W2 [H2][H1] then b2 [H2]  ->  tile entry (k, m), m = 15 for the bias
      const int k = e < K::OFF_B2 ? (e - K::OFF_W2) / H1 : e - K::OFF_B2;
      const int m = e < K::OFF_B2 ? (e - K::OFF_W2) % H1 : 15;
      const int ln = (k >> 2) * 4 + (m >> 2), el = (k & 3) * 4 + (m & 3);
      s = G[(K::G_W2 + el) * 32 + ln] + G[(K::G_W2 + el) * 32 + ln + 16];
    } else {  // W3 [2][H2] then b3 [2]: row 0 = +, row 1 = - of the collapsed row
      const int r = e < K::OFF_B3 ? (e - K::OFF_W3) / H2 : e - K::OFF_B3;
      const int idx = e < K::OFF_B3 ? K::G_W3 + (e - K::OFF_W3) % H2 : K::G_B3;
      for (int l = 0; l < 32; ++l) s += G[idx * 32 + l];
      if (r == 1) s = -s;
    }
    part[3 + e] = s;
  }
}

// ---- host side --------------------------------------------------------------------------------------------------
// Warps per CTA (one CTA per SM): 12 = three per scheduler at <= 168 registers, 8 = two per scheduler at <= 255.
// GPFLOW_SF2_WARPS=8|12 overrides the default for experiments (read once).
inline int sf2_warps() {
  static const int w = [] {
    const char* e = getenv("GPFLOW_SF2_WARPS");
    const int v = e ? atoi(e) : 0;
    return (v == 8 || v == 12) ? v : 12;
  }();
  return w;
}
constexpr int kSF2MaxWarps = 12;

template <int H1, int H2>
size_t sf2_workspace(const GpFlowDesc* d, int64_t B) {
  typedef SF2<H1, H2, kSF2MaxWarps> K;
  // slots = resident warps of the launch (partials / second-level accumulators exist for all of them); only the
  // first min(slots, tiles) slots ever own a tile (warp-major numbering), so only they need a stash region
  const int64_t tiles = (B + 31) / 32;
  const int64_t ctas = tiles < kMaxSMs ? (tiles < 1 ? 1 : tiles) : kMaxSMs;
  const int64_t slots = ctas * kSF2MaxWarps;
  const int64_t stash_slots = tiles < slots ? tiles : slots;
  size_t bytes = 640 * 4;  // packed weights (staging for the constant bank)
  bytes += align_up((size_t)slots * (3 + K::NPARAM) * 4, 256);
  bytes += align_up((size_t)slots * K::NACC * 32 * 4, 256);
  bytes += (size_t)stash_slots * 4 * d->n_steps * K::STASH_BYTES;
  return bytes;
}

template <int H1, int H2, int NW>
int launch_sf2_nw(const GpFlowDesc* d, const Ptrs<float>& p, int64_t B, void* ws, size_t ws_bytes, cudaStream_t st,
                  int sm_count) {
  typedef SF2<H1, H2, NW> K;
  auto kern = gpflow_sf2_kernel<H1, H2, NW, true>;
  const size_t smem = K::smem_bytes(d->n_steps);
  if (smem > 227 * 1024) return GPFLOW_ERR_UNSUPPORTED;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
    cudaGetLastError();
    return GPFLOW_ERR_CUDA;
  }
  const size_t need = sf2_workspace<H1, H2>(d, B);
  if (ws == nullptr || ws_bytes < need) return GPFLOW_ERR_WORKSPACE;
  if (reinterpret_cast<uintptr_t>(ws) & 255) return GPFLOW_ERR_MISALIGNED;
  const int64_t tiles = (B + 31) / 32;
  int64_t grid = tiles;
  if (grid > sm_count) grid = sm_count;
  if (grid < 1) grid = 1;
  const int64_t slots = grid * NW;
  const int npart = 3 + K::NPARAM;

  SF2Args a;
  memset(&a, 0, sizeof(a));
  a.params = p.params; a.x_in = p.x_in; a.x_data = p.x_data; a.grad_y = p.grad_y;
  a.y_out = p.y_out; a.grad_x_out = p.grad_x_out;
  char* wsp = reinterpret_cast<char*>(ws);
  float* packed = reinterpret_cast<float*>(wsp);
  wsp += 640 * 4;
  a.partials = reinterpret_cast<float*>(wsp);
  wsp += align_up((size_t)slots * npart * 4, 256);
  a.g2 = reinterpret_cast<float*>(wsp);
  wsp += align_up((size_t)slots * K::NACC * 32 * 4, 256);
  a.stash = wsp;
  a.B = B; a.n_steps = d->n_steps; a.npart = npart; a.flags = d->flags;
  const bool rev = (d->flags & GPFLOW_FLAG_REVERSE) != 0;
  a.dt = rev ? -(float)d->dt : (float)d->dt;
  a.t0 = rev ? (float)(d->dt * (double)d->n_steps_nominal) : 0.0f;  // res_net.py:217-219
  a.sigma = (float)d->sigma;
  a.loss_scale = (float)d->loss_scale;
#if GPF_SF2_CONST
  sf2_pack_kernel<K><<<5, 128, 0, st>>>(p.params, packed);
  if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  if (cudaMemcpyToSymbolAsync(sf2_cw, packed, 640 * 4, 0, cudaMemcpyDeviceToDevice, st) != cudaSuccess) {
    cudaGetLastError();
    return GPFLOW_ERR_CUDA;
  }
#else
  (void)packed;
#endif
  kern<<<(unsigned)grid, NW * 32, smem, st>>>(a);
  if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  gpflow_finalize_kernel<float><<<(npart + 127) / 128, 128, 0, st>>>(a.partials, (int)slots, npart, npart, p.scalars_out,
                                                                     p.grads_out);
  if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  return GPFLOW_OK;
}

template <int H1, int H2>
int launch_sf2(const GpFlowDesc* d, int op, const Ptrs<float>& p, int64_t B, void* ws, size_t ws_bytes, cudaStream_t st,
               int sm_count) {
  if (op == GPFLOW_OP_FORWARD) return GPFLOW_ERR_UNSUPPORTED;
  if (sf2_warps() == 8) return launch_sf2_nw<H1, H2, 8>(d, p, B, ws, ws_bytes, st, sm_count);
  return launch_sf2_nw<H1, H2, 12>(d, p, B, ws, ws_bytes, st, sm_count);
}

}  // namespace gpflow
