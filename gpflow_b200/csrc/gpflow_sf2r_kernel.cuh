// gpflow_b200 -- sf2r kernel proper + host launch (train step / backward op).  See gpflow_sf2r.cuh for the design; the
// persistent-grid structure, the RK4 loops and the two-level accumulation are those of gpflow_sf2_kernel.cuh.
#pragma once
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "gpflow_dispatch.h"
#include "gpflow_sf2_kernel.cuh"  // SF2Args, kSF2MaxWarps, the first formulation (kept: GPFLOW_SF2_FORM=0)
#include "gpflow_sf2r_rev.cuh"

namespace gpflow {

template <class K>
__global__ void sf2r_pack_kernel(const float* __restrict__ params, float* __restrict__ packed) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < K::WPACK) packed[i] = i < K::WEND ? sf2r_pack_value<K>(params, i) : 0.0f;
}

// gradient entry e of the flat parameter vector from a warp's second-level accumulators G[NACC][32]
template <class K>
GPF_DEV float sf2r_partial_entry(const float* G, int e, const float* G2 = nullptr) {
  typedef typename K::Flat F;
  constexpr int H1 = K::H1, H2 = K::H2;
  float s = 0.0f;
  if (e < F::OFF_W2) {  // W1 [H1][3] then b1 [H1]  ->  tile entry (m & 3, c) of the lanes (slice, ti = m >> 2)
    const int m = e < F::OFF_B1 ? e / 3 : e - F::OFF_B1, c = e < F::OFF_B1 ? e % 3 : 3;
    const int pm = m + 1;   // rows of the dW1 tile are positions
    for (int sl = 0; sl < 8; ++sl) s += G[(K::G_W1 + (pm & 3) * 4 + c) * 32 + sl * 4 + (pm >> 2)];
  } else if (e < F::OFF_W3) {  // W2 [H2][H1] then b2 [H2]  ->  tile entry (position k + 1, position m + 1 | 0 = bias)
    const int pk = (e < F::OFF_B2 ? (e - F::OFF_W2) / H1 : e - F::OFF_B2) + 1;
    const int pm = e < F::OFF_B2 ? (e - F::OFF_W2) % H1 + 1 : 0;
#if GPF_SF2R_MMA
    // C fragments of mma m16n8: lane (g = pk & 7, t = (pm & 7) >> 1), register 4 (pm >> 3) + 2 (pk >> 3) + (pm & 1)
    s = G[(K::G_W2 + 4 * (pm >> 3) + 2 * (pk >> 3) + (pm & 1)) * 32 + (pk & 7) * 4 + ((pm & 7) >> 1)];
#else
    const int ln = (pk >> 2) * 4 + (pm >> 2), el = (pk & 3) * 4 + (pm & 3);
    s = G[(K::G_W2 + el) * 32 + ln] + G[(K::G_W2 + el) * 32 + ln + 16];
#endif
    if (G2 != nullptr) {   // sf2e: the dual evaluations' tile (4 x 4 register tiles of the lanes (slice, ti, tj))
      const int ln2 = (pk >> 2) * 4 + (pm >> 2), el2 = (pk & 3) * 4 + (pm & 3);
      s += G2[el2 * 32 + ln2] + G2[el2 * 32 + ln2 + 16];
    }
  } else if (e < F::OFF_B3) {  // W3 [2][H2]: row 0 = +, row 1 = - of the collapsed row, position k + 1
    const int r = (e - F::OFF_W3) / H2, p = (e - F::OFF_W3) % H2 + 1;
    for (int sl = 0; sl < 8; ++sl) s += G[(K::G_W3 + (p & 3)) * 32 + sl * 4 + (p >> 2)];
    if (r == 1) s = -s;
  } else {  // b3 [2]
    for (int l = 0; l < 32; ++l) s += G[K::G_B3 * 32 + l];
    if (e - F::OFF_B3 == 1) s = -s;
  }
  return s;
}

template <int H1, int H2, int NW>
__global__ void __launch_bounds__(NW * 32, 1) gpflow_sf2r_kernel(const SF2Args a) {
  typedef SF2R<H1, H2, NW> K;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint64_t bar_stage;
  __shared__ uint64_t bars[NW][2];
  float* const W = reinterpret_cast<float*>(smem_raw);
  float* const sct = W + K::WEND;  // [4][4] stage constants
  float* const times = sct + K::SCT;
  float* const raw = times + round_up(a.n_steps + 1, 4);
  float* const warp_area = raw + round_up(K::NPARAM, 4);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, l7 = lane & 7;

  // ---- prologue: TMA bulk copy of the flat parameters, re-pack, stage constants, time table ----
  stage_params_to_smem<float>(raw, a.params, K::NPARAM, &bar_stage);
  for (int i = tid; i < K::WEND; i += NW * 32) W[i] = sf2r_pack_value<K>(raw, i);
  if (tid < 4) {
    // RK4 stage sg (ode_order_method.py:66-71): time offset, weight of k_sg in the sum, step to the next stage point
    // (0 after the last stage; also the adjoint's dt-factor of yb_{sg+1}), weight * dt/6
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
  float* const Sbuf0 = wa;                 // stash buffer b = Sbuf0 + b * S_FL
  float* const Wk = wa + 2 * K::S_FL;      // work rows of the reverse sweep
  // warp-major slot numbering: a small batch (tiles < resident warps) spreads over the SMs first
  const int slot = warp * gridDim.x + blockIdx.x, nslots = gridDim.x * NW;
  float* const G = a.g2 + (size_t)slot * K::NACC * 32;
  for (int i = 0; i < K::NACC; ++i) G[i * 32 + lane] = 0.0f;
  __syncthreads();

  const bool wrap = (a.flags & 0x1u) != 0;
  const float bnd = 1.0f - Eps<float>::value;  // velocity_field.py:96
  const float c6 = a.dt / 6.0f;
  const float rsq2s = 1.0f / (1.4142135623730951f * a.sigma), sq2s = 1.4142135623730951f * a.sigma;
  const int E = 4 * a.n_steps;
  char* const stash = a.stash + (size_t)slot * E * K::STASH_BYTES;
  uint32_t phbits = 0;  // bit b = mbarrier phase parity of stash buffer b
  uint64_t* const mybars = &bars[warp][0];

  float acc[16], accw[16], acc3[4], b3b = 0.0f;
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = accw[i] = 0.0f;
#pragma unroll
  for (int i = 0; i < 4; ++i) acc3[i] = 0.0f;

  const long long ntiles = (a.B + 31) / 32;
  if (a.stagger_cycles != 0) {
    // > 0: odd warps start `stagger_cycles` late; < 0: warp w starts (w * 5 mod NW) / NW * |stagger_cycles| late
    const long long wait = a.stagger_cycles > 0 ? ((warp & 1) ? a.stagger_cycles : 0)
                                                : (long long)(-a.stagger_cycles) * ((warp * 5) % NW) / NW;
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
    // optional phase alignment of the CTA's warps, one named barrier per RK4 step (see gpflow_sf2e_kernel, where it is worth
    // +36 %; measured here: see launch_sf2r)
    long long n_act = (ntiles - (tile - slot) - blockIdx.x + gridDim.x - 1) / gridDim.x;
    if (n_act > NW) n_act = NW;
    const unsigned n_act_threads = 32u * (unsigned)n_act;
    auto cta_align = [&]() {
      if (a.cta_align && n_act_threads > 32u) asm volatile("bar.sync 1, %0;" ::"r"(n_act_threads) : "memory");
    };

    // ---------------- forward sweep: 4N velocity evaluations, classical RK4 (ode_order_method.py:66-71) ----------
    float ys0 = x0, ys1 = x1, ks0 = 0.f, ks1 = 0.f;
#pragma unroll 1
    for (int e = 0; e < E; ++e) {
      const int sg = e & 3, buf = e & 1;
      if (sg == 0) cta_align();
      const float4 sc = ldsw4(sct + 4 * sg);
      const float ts = times[e >> 2] + sc.x;  // t, t + dt/2, t + dt/2, t + dt
      float* const Sb = Sbuf0 + buf * K::S_FL;
      if (e >= 2) {  // the bulk store that read this buffer two evaluations ago must have drained
        if (lane == 0) bulk_wait_read<1>();
        __syncwarp();
      }
      float v0, v1;
      sf2r_forward_eval<K>(W, Sb + lane * K::SROW, l7, ys0, ys1, ts, bnd, v0, v1);
      fence_proxy_async();  // generic-proxy row writes -> visible to the async proxy (TMA)
      __syncwarp();
      if (lane == 0) {
        bulk_s2g(stash + (size_t)e * K::STASH_BYTES, Sb, K::STASH_BYTES);
        bulk_commit();
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
      bulk_g2s(Sbuf0 + ((E - 1) & 1) * K::S_FL, stash + (size_t)(E - 1) * K::STASH_BYTES, K::STASH_BYTES,
               mybars + ((E - 1) & 1));
    }
    // ---------------- reverse sweep (A.5) ----------------
    float sum0 = 0.f, sum1 = 0.f, prev0 = 0.f, prev1 = 0.f;
#pragma unroll 1
    for (int e = E - 1; e >= 0; --e) {
      const int sg = e & 3, buf = e & 1;
      if (sg == 3) cta_align();
      const float4 sc = ldsw4(sct + 4 * sg);
      const float ts = times[e >> 2] + sc.x;
      // prefetch the rows of evaluation e-1 into the other buffer (its previous user, e+1, is complete); the same
      // barrier orders this evaluation's work-row writes after the previous evaluation's last reads
      fence_proxy_async();
      __syncwarp();
      if (e > 0 && lane == 0) {
        const int nb = buf ^ 1;
        mbar_expect_tx(mybars + nb, K::STASH_BYTES);
        bulk_g2s(Sbuf0 + nb * K::S_FL, stash + (size_t)(e - 1) * K::STASH_BYTES, K::STASH_BYTES, mybars + nb);
      }
      mbar_wait(mybars + buf, (phbits >> buf) & 1u);
      phbits ^= 1u << buf;
      // kb_4 = dt/6 xb; kb_3 = dt/3 xb + dt yb_4; kb_2 = dt/3 xb + dt/2 yb_3; kb_1 = dt/6 xb + dt/2 yb_2
      const float kb0 = fmaf(sc.z, prev0, sc.w * xb0), kb1 = fmaf(sc.z, prev1, sc.w * xb1);
      float r0, r1;
      sf2r_reverse_eval<K>(W, Sbuf0 + buf * K::S_FL, Wk, lane, ts, bnd, kb0, kb1, r0, r1, acc, accw, acc3, b3b);
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
      G[(K::G_W1 + i) * 32 + lane] += accw[i];
      acc[i] = accw[i] = 0.0f;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      G[(K::G_W3 + i) * 32 + lane] += acc3[i];
      acc3[i] = 0.0f;
    }
    G[K::G_B3 * 32 + lane] += b3b;
    b3b = 0.0f;
    G[K::G_OT * 32 + lane] += ot;
  }

  // ---------------- per-warp partial vector [ot, reg, lagr | grads] ----------------
  __syncwarp();
  float* part = a.partials + (size_t)slot * a.npart;
  const float ot_w = warp_sum(G[K::G_OT * 32 + lane]);
  if (lane == 0) {
    part[0] = ot_w * a.loss_scale;
    part[1] = 0.0f;
    part[2] = 0.0f;
  }
  for (int e = lane; e < K::NPARAM; e += 32) part[3 + e] = sf2r_partial_entry<K>(G, e);
}

// ---- host side --------------------------------------------------------------------------------------------------
template <int H1, int H2>
size_t sf2r_workspace(const GpFlowDesc* d, int64_t B) {
  typedef SF2R<H1, H2, kSF2MaxWarps> K;
  const int64_t tiles = (B + 31) / 32;
  const int64_t ctas = tiles < kMaxSMs ? (tiles < 1 ? 1 : tiles) : kMaxSMs;
  const int64_t slots = ctas * kSF2MaxWarps;
  const int64_t stash_slots = tiles < slots ? tiles : slots;
  size_t bytes = 704 * 4;  // packed weights (staging for the constant bank)
  bytes += align_up((size_t)slots * (3 + K::NPARAM) * 4, 256);
  bytes += align_up((size_t)slots * K::NACC * 32 * 4, 256);
  bytes += (size_t)stash_slots * 4 * d->n_steps * K::STASH_BYTES;
  return bytes;
}

// The constant-bank copy of the packed weights (sf2r_cw) is ONE array per device, rewritten by every call.  Calls on one
// stream are ordered by the stream; a call on another stream must not overwrite the array while the previous call's
// kernel is still reading it, so every call waits (on the device, cudaStreamWaitEvent) for the event recorded behind the
// previous call's kernel on this device.  The mutex keeps two host threads from interleaving their [wait, copy, launch,
// record] sequences.  The event is created on the first call per device (not device memory).  While the stream is being
// captured into a CUDA graph the wait / record are skipped: the captured nodes are ordered by the capture itself, and
// replays of such a graph must not overlap other calls of this kernel on the same device.
struct SF2RConstGuard {
  std::mutex mu;
  cudaEvent_t ev[64];
  bool have[64];
};
inline SF2RConstGuard& sf2r_const_guard() {
  static SF2RConstGuard g = {};
  return g;
}

// stage the packed weights into the constant bank and launch the main kernel (`launch()`), ordered as described above
template <class K, class F>
int sf2r_launch_guarded(const float* params, float* packed, cudaStream_t st, F launch) {
#if GPF_SF2R_CONST
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
    cudaGetLastError();
    return GPFLOW_ERR_CUDA;
  }
  SF2RConstGuard& g = sf2r_const_guard();
  std::lock_guard<std::mutex> lk(g.mu);
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cs) != cudaSuccess) {
    cudaGetLastError();
    cs = cudaStreamCaptureStatusNone;
  }
  const bool guard = cs == cudaStreamCaptureStatusNone;
  if (guard) {
    if (!g.have[dev]) {
      if (cudaEventCreateWithFlags(&g.ev[dev], cudaEventDisableTiming) != cudaSuccess) {
        cudaGetLastError();
        return GPFLOW_ERR_CUDA;
      }
      g.have[dev] = true;
    } else if (cudaStreamWaitEvent(st, g.ev[dev], 0) != cudaSuccess) {
      cudaGetLastError();
      return GPFLOW_ERR_CUDA;
    }
  }
  sf2r_pack_kernel<K><<<(K::WPACK + 127) / 128, 128, 0, st>>>(params, packed);
  if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  if (cudaMemcpyToSymbolAsync(sf2r_cw, packed, K::WPACK * 4, 0, cudaMemcpyDeviceToDevice, st) != cudaSuccess) {
    cudaGetLastError();
    return GPFLOW_ERR_CUDA;
  }
  launch();
  if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  if (guard && cudaEventRecord(g.ev[dev], st) != cudaSuccess) {
    cudaGetLastError();
    return GPFLOW_ERR_CUDA;
  }
  return GPFLOW_OK;
#else
  (void)params;
  (void)packed;
  launch();
  return cudaGetLastError() == cudaSuccess ? GPFLOW_OK : GPFLOW_ERR_CUDA;
#endif
}

template <int H1, int H2>
int launch_sf2r(const GpFlowDesc* d, const Ptrs<float>& p, int64_t B, void* ws, size_t ws_bytes, cudaStream_t st,
                int sm_count) {
  constexpr int NW = kSF2MaxWarps;
  typedef SF2R<H1, H2, NW> K;
  auto kern = gpflow_sf2r_kernel<H1, H2, NW>;
  const size_t smem = K::smem_bytes(d->n_steps);
  if (smem > 227 * 1024) return GPFLOW_ERR_UNSUPPORTED;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
    cudaGetLastError();
    return GPFLOW_ERR_CUDA;
  }
  if (ws == nullptr || ws_bytes < sf2r_workspace<H1, H2>(d, B)) return GPFLOW_ERR_WORKSPACE;
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
  wsp += 704 * 4;
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
  // All warps of an SM start their first tile together and would run the FMA-heavy forward sweeps and the shared-memory-
  // heavy reverse sweeps in phase; spreading their start over ~half a tile time (one tile = 120 evaluations, ~50 k cycles
  // per RK4 step) was measured at -2.4 % on a 2^21-sample launch and -0.3 % at 2^24 (GPFLOW_SF2R_STAGGER = cycles,
  // 0 = off; negative = spread over the warps, positive = odd warps late).  Only when every warp has >= 4 tiles.
  static const int stagger_env = [] { const char* e = getenv("GPFLOW_SF2R_STAGGER"); return e ? atoi(e) : 1; }();
  a.stagger_cycles = stagger_env != 1 ? stagger_env : (tiles >= 4 * slots ? -26000 * d->n_steps : 0);
  // GPFLOW_SF2R_ALIGN=1: one named barrier per RK4 step over the CTA's warps.  Worth +36 % in gpflow_sf2e_kernel (instruction
  // cache); measured here on the C4 headline, alternating: 108.6 / 108.8 ms free-running, 124.5 / 124.5 ms aligned -- this
  // kernel fits the instruction cache and is bound by the FMA and shared-memory pipes, which aligned warps hit in phase.
  static const int align_env = [] { const char* e = getenv("GPFLOW_SF2R_ALIGN"); return e ? atoi(e) : 0; }();
  a.cta_align = align_env;
  if (a.cta_align) a.stagger_cycles = 0;   // aligned warps start together
  const int rc = sf2r_launch_guarded<K>(p.params, packed, st, [&]() { kern<<<(unsigned)grid, NW * 32, smem, st>>>(a); });
  if (rc != GPFLOW_OK) return rc;
  gpflow_finalize_wide_kernel<float><<<(npart + 31) / 32, 256, 0, st>>>(a.partials, (int)slots, npart, npart, p.scalars_out,
                                                                        p.grads_out);
  if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  return GPFLOW_OK;
}

}  // namespace gpflow
