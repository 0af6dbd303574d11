// gpflow_b200 -- host-side launch of the small-d kernels (included by the inst_small_*.cu translation units)
#pragma once
#include <string.h>

#include "gpflow_dispatch.h"
#include "gpflow_small.cuh"

namespace gpflow {

template <typename T, int D, int NB, int H1, int H2, int MODE, bool EULER>
int launch_small_variant(const GpFlowDesc* d, const Ptrs<T>& p, int64_t B, void* ws, size_t ws_bytes, cudaStream_t st,
                         int sm_count) {
  typedef SmallKernel<T, D, NB, H1, H2, kSmallBlock> K;
  auto kern = gpflow_small_kernel<T, D, NB, H1, H2, kSmallBlock, MODE, EULER>;
  const size_t smem = K::smem_bytes(MODE, d->n_steps);
  if (smem > 227 * 1024) return GPFLOW_ERR_UNSUPPORTED;
  // kernel attributes: idempotent and cheap; set on every call so the library stays stateless across devices
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
    cudaGetLastError();
    return GPFLOW_ERR_CUDA;
  }
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kSmallBlock, smem) != cudaSuccess || occ < 1) {
    cudaGetLastError();
    return GPFLOW_ERR_CUDA;
  }
  int64_t grid = (B + kSmallBlock - 1) / kSmallBlock;
  const int64_t resident = (int64_t)sm_count * occ;
  if (grid > resident) grid = resident;
  if (grid * kSmallBlock > (int64_t)kMaxSMs * 2048) grid = (int64_t)kMaxSMs * 2048 / kSmallBlock;
  const int npart = 3 + K::NPARAM;
  const size_t need = small_workspace(d, B, MODE == MODE_FWD ? GPFLOW_OP_FORWARD : GPFLOW_OP_TRAIN_STEP, K::NPARAM);
  if (ws == nullptr || ws_bytes < need) return GPFLOW_ERR_WORKSPACE;
  if (reinterpret_cast<uintptr_t>(ws) & 15) return GPFLOW_ERR_MISALIGNED;

  KArgs<T> a;
  memset(&a, 0, sizeof(a));
  a.params = p.params; a.x_in = p.x_in; a.x_data = p.x_data; a.grad_y = p.grad_y; a.reg_cotangent = p.reg_cot;
  a.zw = p.zw; a.y_out = p.y_out; a.logdet_out = p.logdet_out; a.traj_out = p.traj_out; a.grad_x_out = p.grad_x_out;
  char* wsp = reinterpret_cast<char*>(ws);
  a.partials = reinterpret_cast<T*>(wsp);
  const int64_t max_threads = (int64_t)kMaxSMs * 2048;
  int64_t ts_max = (int64_t)align_up((size_t)B, kSmallBlock);
  if (ts_max > max_threads) ts_max = max_threads;
  wsp += align_up((size_t)(ts_max / kSmallBlock) * npart * sizeof(T), 256);
  a.traj_ws = reinterpret_cast<T*>(wsp);
  a.B = B; a.n_steps = d->n_steps; a.npart = npart; a.flags = d->flags;
  const bool rev = (d->flags & GPFLOW_FLAG_REVERSE) != 0;
  a.dt = rev ? -(T)d->dt : (T)d->dt;
  a.t0 = rev ? (T)(d->dt * (double)d->n_steps_nominal) : (T)0;  // res_net.py:217-219
  a.dt_constraint = (T)d->dt_constraint; a.sigma = (T)d->sigma; a.euler_coef = (T)d->euler_coef;
  a.loss_scale = (T)d->loss_scale;
  kern<<<(unsigned)grid, kSmallBlock, smem, st>>>(a);
  if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  const int count = MODE == MODE_FWD ? 3 : npart;
  if (p.scalars_out != nullptr || p.grads_out != nullptr) {
    gpflow_finalize_kernel<T><<<(count + 127) / 128, 128, 0, st>>>(a.partials, (int)grid, npart, count, p.scalars_out,
                                                                   p.grads_out);
    if (cudaGetLastError() != cudaSuccess) return GPFLOW_ERR_CUDA;
  }
  return GPFLOW_OK;
}

template <typename T, int D, int NB, int H1, int H2, bool EULER>
int launch_small_shape(const GpFlowDesc* d, int op, const void* ptrs, int64_t B, void* ws, size_t wsb, cudaStream_t st,
                       int sm_count) {
  const Ptrs<T>& p = *reinterpret_cast<const Ptrs<T>*>(ptrs);
  if (op == GPFLOW_OP_FORWARD) return launch_small_variant<T, D, NB, H1, H2, MODE_FWD, EULER>(d, p, B, ws, wsb, st, sm_count);
  if (op == GPFLOW_OP_TRAIN_STEP) return launch_small_variant<T, D, NB, H1, H2, MODE_TRAIN, EULER>(d, p, B, ws, wsb, st, sm_count);
  return launch_small_variant<T, D, NB, H1, H2, MODE_BWD, EULER>(d, p, B, ws, wsb, st, sm_count);
}

}  // namespace gpflow
