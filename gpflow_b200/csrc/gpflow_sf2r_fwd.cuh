// gpflow_b200 -- sf2r forward-only kernel (the round-2 evaluation: grad psi in reverse mode, two contractions instead of
// three, weights from the constant bank, no per-lane shared-memory rows at all): transport (forward or reverse in time)
// of the 2-D stream-function flow, fp32.  The evaluation path behind `compute_ot_costs` / `ot_map` /
// `gp_flow(reverse=True)` (src/plots/outputs.py:7-248,
// network.py:156-253): y, log|det|, optional trajectory (ode_solver.py:70-81) and the Lagrangian diagnostic
// (res_net.py:292-294, pairs x_{i+1} with t_i).  Same per-evaluation code as the train kernel (sf2r_forward_eval without the
// stash stores): registers only, 16 warps per SM.  gpflow_sf2_fwd.cuh is the round-1 version (GPFLOW_SF2_FORM=0).
#pragma once
#include "gpflow_sf2_fwd.cuh"
#include "gpflow_sf2r_kernel.cuh"

namespace gpflow {

template <int H1, int H2, int NW>
__global__ void __launch_bounds__(NW * 32, 1) gpflow_sf2r_fwd_kernel(const SF2FwdArgs a) {
  typedef SF2R<H1, H2, NW> K;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ uint64_t bar_stage;
  float* const W = reinterpret_cast<float*>(smem_raw);
  float* const sct = W + K::WEND;
  float* const times = sct + K::SCT;
  float* const raw = times + round_up(a.n_steps + 1, 4);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  stage_params_to_smem<float>(raw, a.params, K::NPARAM, &bar_stage);
  for (int i = tid; i < K::WEND; i += NW * 32) W[i] = sf2r_pack_value<K>(raw, i);
  if (tid < 4) {
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
  __syncthreads();

  const bool wrap = (a.flags & 0x1u) != 0, want_lagr = (a.flags & 0x2u) != 0;
  const float bnd = 1.0f - Eps<float>::value;
  const float c6 = a.dt / 6.0f;
  const float rsq2s = 1.0f / (1.4142135623730951f * a.sigma), sq2s = 1.4142135623730951f * a.sigma;
  const int nsg = want_lagr ? 5 : 4;  // stage 4 = the Lagrangian's extra velocity evaluation at (x_{i+1}, t_i)
  // warp-major slot numbering: a small batch (tiles < resident warps) spreads over the SMs first, then over the
  // warps of an SM
  const int slot = warp * gridDim.x + blockIdx.x, nslots = gridDim.x * NW;
  float lagr = 0.0f;

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
    float ys0 = x0, ys1 = x1, ks0 = 0.f, ks1 = 0.f;
    int step = 0, sg = 0;
#pragma unroll 1
    for (int e = 0; e < nsg * a.n_steps; ++e) {
      if (sg == 0 && a.traj_out != nullptr && valid) {
        a.traj_out[((long long)step * a.B + b) * 2] = x0;
        a.traj_out[((long long)step * a.B + b) * 2 + 1] = x1;
      }
      const float4 sc = ldsw4(sct + 4 * (sg & 3));
      const float ts = sg == 4 ? times[step] : times[step] + sc.x;
      float v0, v1;
      sf2r_forward_eval<K, false>(W, nullptr, 0, ys0, ys1, ts, bnd, v0, v1);
      if (sg == 4) {
        if (valid) lagr += 0.5f * fmaf(v0, v0, v1 * v1);
      } else {
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
      if (++sg == nsg) {
        sg = 0;
        ++step;
      }
    }
    if (a.traj_out != nullptr && valid) {
      a.traj_out[((long long)a.n_steps * a.B + b) * 2] = x0;
      a.traj_out[((long long)a.n_steps * a.B + b) * 2 + 1] = x1;
    }
    float y0 = x0, y1 = x1;
    if (wrap) {  // network.py:178-185
      const float lim = 1.0f - Eps<float>::value;
      y0 = sq2s * erfinv_real(fminf(fmaxf(x0, -lim), lim));
      y1 = sq2s * erfinv_real(fminf(fmaxf(x1, -lim), lim));
    }
    if (valid) {
      a.y_out[2 * b] = y0;
      a.y_out[2 * b + 1] = y1;
      if (a.logdet_out != nullptr)
        a.logdet_out[b] = (fmaf(y0, y0, y1 * y1) - fmaf(xin0, xin0, xin1 * xin1)) / (2.0f * a.sigma * a.sigma);
    }
  }
  lagr = warp_sum(lagr);
  if (lane == 0) {
    float* part = a.partials + (size_t)slot * 3;
    part[0] = 0.0f;
    part[1] = 0.0f;
    part[2] = lagr * a.dt_constraint * a.loss_scale;
  }
}

template <int H1, int H2>
size_t sf2r_fwd_workspace(int64_t B) {
  (void)B;
  return 704 * 4 + align_up((size_t)kMaxSMs * kSF2FwdWarps * 3 * 4, 256);
}

template <int H1, int H2>
int launch_sf2r_fwd(const GpFlowDesc* d, const Ptrs<float>& p, int64_t B, void* ws, size_t ws_bytes, cudaStream_t st,
                   int sm_count) {
  constexpr int NW = kSF2FwdWarps;
  typedef SF2R<H1, H2, NW> K;
  auto kern = gpflow_sf2r_fwd_kernel<H1, H2, NW>;
  const size_t smem = (size_t)(K::WEND + K::SCT + round_up(d->n_steps + 1, 4) + round_up(K::NPARAM, 4)) * sizeof(float) + 128;
  if (smem > 227 * 1024) return GPFLOW_ERR_UNSUPPORTED;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
    cudaGetLastError();
    return GPFLOW_ERR_CUDA;
  }
  if (ws == nullptr || ws_bytes < sf2r_fwd_workspace<H1, H2>(B)) return GPFLOW_ERR_WORKSPACE;
  if (reinterpret_cast<uintptr_t>(ws) & 15) return GPFLOW_ERR_MISALIGNED;
  const int64_t tiles = (B + 31) / 32;
  int64_t grid = tiles;
  if (grid > sm_count) grid = sm_count;
  if (grid < 1) grid = 1;
  SF2FwdArgs a;
  memset(&a, 0, sizeof(a));
  a.params = p.params; a.x_in = p.x_in; a.y_out = p.y_out; a.logdet_out = p.logdet_out; a.traj_out = p.traj_out;
  float* packed = reinterpret_cast<float*>(ws);
  a.partials = packed + 704;
  a.B = B; a.n_steps = d->n_steps; a.flags = d->flags;
  const bool rev = (d->flags & GPFLOW_FLAG_REVERSE) != 0;
  a.dt = rev ? -(float)d->dt : (float)d->dt;
  a.t0 = rev ? (float)(d->dt * (double)d->n_steps_nominal) : 0.0f;  // res_net.py:217-219
  a.sigma = (float)d->sigma; a.dt_constraint = (float)d->dt_constraint; a.loss_scale = (float)d->loss_scale;
  const int rc = sf2r_launch_guarded<K>(p.params, packed, st, [&]() { kern<<<(unsigned)grid, NW * 32, smem, st>>>(a); });
  if (rc != GPFLOW_OK) return rc;
  if (p.scalars_out != nullptr) {
    gpflow_finalize_kernel<float><<<1, 128, 0, st>>>(a.partials, (int)(grid * NW), 3, 3, p.scalars_out, nullptr);
    if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  }
  return GPFLOW_OK;
}

}  // namespace gpflow
