// gpflow_b200 -- sf2 kernel proper + host launch.  See gpflow_sf2.cuh for the design.
#pragma once
#include <string.h>

#include "gpflow_dispatch.h"
#include "gpflow_sf2_eval.cuh"

namespace gpflow {

template <int H1, int H2, int NW, bool TRAIN>
__global__ void __launch_bounds__(NW * 32, 1) gpflow_sf2_kernel(const SF2Args a) {
  typedef SF2<H1, H2, NW> K;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t bar_stage;
  __shared__ uint64_t bars[NW][2];
  float* const W = reinterpret_cast<float*>(smem_raw);
  float* const times = W + K::WEND;
  float* const raw = times + round_up(a.n_steps + 1, 4);
  float* const warp_area = raw + round_up(K::NPARAM, 4);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // ---- prologue: TMA bulk copy of the flat parameters, re-pack (transpose / collapse / pad), time table ----
  stage_params_to_smem<float>(raw, a.params, K::NPARAM, &bar_stage);
  for (int i = tid; i < K::WEND; i += NW * 32) {
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
    W[i] = v;
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
  float* const Cbuf = wa + 2 * K::A_FL + 2 * K::B_FL;
  float* const Crow = Cbuf + lane * K::CS;
  const int slot = blockIdx.x * NW + warp, nslots = gridDim.x * NW;
  float* const G = a.g2 + (size_t)slot * K::NACC * 32;
  if (TRAIN) {
#pragma unroll
    for (int i = 0; i < K::CS / 4; ++i) sts4(Crow + 4 * i, 0.f, 0.f, 0.f, 0.f);
    for (int i = 0; i < K::NACC; ++i) G[i * 32 + lane] = 0.0f;
  }
  __syncthreads();

  const bool wrap = (a.flags & 0x1u) != 0;
  const float bnd = 1.0f - Eps<float>::value;  // velocity_field.py:96
  const float dt = a.dt, hdt = a.dt / 2.0f, c6 = a.dt / 6.0f;
  const float rsq2s = 1.0f / (1.4142135623730951f * a.sigma), sq2s = 1.4142135623730951f * a.sigma;
  const int E = 4 * a.n_steps;
  char* const stash = a.stash + (size_t)slot * E * K::STASH_BYTES;
  uint32_t ph0 = 0, ph1 = 0;  // mbarrier phase parity of the two stash buffers

  float acc[16], w3b[16], b3b = 0.0f;
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = w3b[i] = 0.0f;

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
      const float t = times[e >> 2];
      const float ts = sg == 0 ? t : (sg == 3 ? t + dt : t + hdt);
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
      const float wk = (sg == 1 || sg == 2) ? 2.0f : 1.0f;
      const float ak = sg == 2 ? dt : hdt;
      ks0 = sg == 0 ? v0 : fmaf(wk, v0, ks0);
      ks1 = sg == 0 ? v1 : fmaf(wk, v1, ks1);
      if (sg == 3) {
        x0 = fmaf(c6, ks0, x0);
        x1 = fmaf(c6, ks1, x1);
        ys0 = x0;
        ys1 = x1;
      } else {
        ys0 = fmaf(ak, v0, x0);
        ys1 = fmaf(ak, v1, x1);
      }
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
      mbar_expect_tx(&bars[warp][(E - 1) & 1], K::STASH_BYTES);
      const char* src = stash + (size_t)(E - 1) * K::STASH_BYTES;
      bulk_g2s(Abuf0 + ((E - 1) & 1) * K::A_FL, src, K::A_BYTES, &bars[warp][(E - 1) & 1]);
      bulk_g2s(Bbuf0 + ((E - 1) & 1) * K::B_FL, src + K::A_BYTES, K::B_BYTES, &bars[warp][(E - 1) & 1]);
    }
    // ---------------- reverse sweep (A.5) ----------------
    float sum0 = 0.f, sum1 = 0.f, prev0 = 0.f, prev1 = 0.f;
#pragma unroll 1
    for (int e = E - 1; e >= 0; --e) {
      const int sg = e & 3, buf = e & 1;
      const float t = times[e >> 2];
      const float ts = sg == 0 ? t : (sg == 3 ? t + dt : t + hdt);
      // prefetch the rows of evaluation e-1 into the other buffer (its previous user, e+1, is complete)
      fence_proxy_async();
      __syncwarp();
      if (e > 0 && lane == 0) {
        const int nb = buf ^ 1;
        mbar_expect_tx(&bars[warp][nb], K::STASH_BYTES);
        const char* src = stash + (size_t)(e - 1) * K::STASH_BYTES;
        bulk_g2s(Abuf0 + nb * K::A_FL, src, K::A_BYTES, &bars[warp][nb]);
        bulk_g2s(Bbuf0 + nb * K::B_FL, src + K::A_BYTES, K::B_BYTES, &bars[warp][nb]);
      }
      if (buf == 0) { mbar_wait(&bars[warp][0], ph0); ph0 ^= 1u; }
      else { mbar_wait(&bars[warp][1], ph1); ph1 ^= 1u; }
      // kb_4 = dt/6 xb; kb_3 = dt/3 xb + dt yb_4; kb_2 = dt/3 xb + dt/2 yb_3; kb_1 = dt/6 xb + dt/2 yb_2
      const float wk = (sg == 1 || sg == 2) ? 2.0f * c6 : c6;
      const float ak = sg == 3 ? 0.0f : (sg == 2 ? dt : hdt);
      const float kb0 = fmaf(ak, prev0, wk * xb0), kb1 = fmaf(ak, prev1, wk * xb1);
      float r0, r1;
      sf2_reverse_eval<K>(W, Abuf0 + buf * K::A_FL, Bbuf0 + buf * K::B_FL, Crow, lane, ts, bnd, kb0, kb1, r0, r1, acc,
                          w3b, b3b);
      prev0 = r0; prev1 = r1;
      sum0 += r0; sum1 += r1;
      if (sg == 0) {
        xb0 += sum0; xb1 += sum1;
        sum0 = sum1 = prev0 = prev1 = 0.0f;
      }
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
      acc[i] = w3b[i] = 0.0f;
    }
    G[K::G_B3 * 32 + lane] += b3b;
    b3b = 0.0f;
    G[K::G_OT * 32 + lane] += ot;
    __syncwarp();
    float c0 = 0.f, c1 = 0.f;
#pragma unroll 8
    for (int r = 0; r < 32; ++r) {
      c0 += Cbuf[r * K::CS + lane];
      c1 += Cbuf[r * K::CS + 32 + lane];
    }
    G[K::G_W1 * 32 + lane] += c0;
    G[(K::G_W1 + 1) * 32 + lane] += c1;
    __syncwarp();
#pragma unroll
    for (int i = 0; i < K::CS / 4; ++i) sts4(Crow + 4 * i, 0.f, 0.f, 0.f, 0.f);
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
    if (e < K::OFF_W2) {  // W1 [H1][3] then b1 [H1]  ->  column 4m + c of the C rows
      const int col = e < K::OFF_B1 ? 4 * (e / 3) + e % 3 : 4 * (e - K::OFF_B1) + 3;
      s = G[(K::G_W1 + (col >> 5)) * 32 + (col & 31)];
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
constexpr int kSF2Warps = 8;

template <int H1, int H2>
size_t sf2_workspace(const GpFlowDesc* d, int64_t B) {
  typedef SF2<H1, H2, kSF2Warps> K;
  const int64_t max_slots = (int64_t)kMaxSMs * kSF2Warps;
  int64_t slots = (B + 31) / 32;
  slots = (slots + kSF2Warps - 1) / kSF2Warps * kSF2Warps;
  if (slots > max_slots) slots = max_slots;
  if (slots < kSF2Warps) slots = kSF2Warps;
  size_t bytes = align_up((size_t)slots * (3 + K::NPARAM) * 4, 256);
  bytes += align_up((size_t)slots * K::NACC * 32 * 4, 256);
  bytes += (size_t)slots * 4 * d->n_steps * K::STASH_BYTES;
  return bytes;
}

template <int H1, int H2>
int launch_sf2(const GpFlowDesc* d, int op, const Ptrs<float>& p, int64_t B, void* ws, size_t ws_bytes, cudaStream_t st,
               int sm_count) {
  typedef SF2<H1, H2, kSF2Warps> K;
  if (op == GPFLOW_OP_FORWARD) return GPFLOW_ERR_UNSUPPORTED;
  auto kern = gpflow_sf2_kernel<H1, H2, kSF2Warps, true>;
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
  int64_t grid = (tiles + kSF2Warps - 1) / kSF2Warps;
  if (grid > sm_count) grid = sm_count;
  if (grid < 1) grid = 1;
  const int64_t slots = grid * kSF2Warps;
  const int npart = 3 + K::NPARAM;

  SF2Args a;
  memset(&a, 0, sizeof(a));
  a.params = p.params; a.x_in = p.x_in; a.x_data = p.x_data; a.grad_y = p.grad_y;
  a.y_out = p.y_out; a.grad_x_out = p.grad_x_out;
  char* wsp = reinterpret_cast<char*>(ws);
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
  kern<<<(unsigned)grid, kSF2Warps * 32, smem, st>>>(a);
  if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  gpflow_finalize_kernel<float><<<(npart + 127) / 128, 128, 0, st>>>(a.partials, (int)slots, npart, npart, p.scalars_out,
                                                                     p.grads_out);
  if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  return GPFLOW_OK;
}

}  // namespace gpflow
