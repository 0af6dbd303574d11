// gpflow_b200 -- gpflow_velocity_static(): the velocity evaluation for callers that cannot provide (or allocate) a
// workspace -- code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in
// the reference (see include/gpflow_b200.h).
//
// The per-warp scratch of the generic velocity kernel (activations and tangents of one evaluation, gpflow_generic.cuh)
// lives in ONE file-scope device array of GPFLOW_STATIC_SCRATCH_BYTES.  Every GPU that loads the library holds that
// array until the process exits.  Every call shares it: calls that overlap on the GPU (different streams, different host
// threads) race on it; calls on one stream are ordered and safe.  The array keeps the last call's contents; nothing
// in the velocity evaluation relies on its initial state (the generic path clears the one region it needs zeroed, the
// padded parameter copy, itself with cudaMemsetAsync on the caller's stream), so nothing is cleared here.
//
// Host code must not hand `g_static_scratch` itself to a kernel: in host code the name is a host-side stand-in, not a
// device address.  The device address is looked up once per device, with that device current, by
// cudaGetSymbolAddress() below and cached.
#include "../../include/gpflow_b200.h"

#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

namespace {

static __device__ __align__(256) unsigned char g_static_scratch[GPFLOW_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;

// device address of g_static_scratch on the CURRENT device (set-up: once per device), nullptr + *rc on failure
void* static_scratch_ptr(int* rc) {
  static std::mutex mu;
  static void* cache[kMaxDevices] = {nullptr};
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    cudaGetLastError();
    *rc = GPFLOW_ERR_CUDA;
    return nullptr;
  }
  std::lock_guard<std::mutex> lk(mu);
  if (cache[dev] == nullptr) {
    void* p = nullptr;
    if (cudaGetSymbolAddress(&p, g_static_scratch) != cudaSuccess || p == nullptr) {
      cudaGetLastError();
      *rc = GPFLOW_ERR_CUDA;
      return nullptr;
    }
    cache[dev] = p;
  }
  *rc = GPFLOW_OK;
  return cache[dev];
}

}  // namespace

extern "C" int gpflow_velocity_static(const GpFlowDesc* desc, int64_t B, const void* params, const void* x, double t,
                                      void* v_out, void* cuda_stream) {
  if (desc == nullptr) return GPFLOW_ERR_NULL_POINTER;
  GpFlowDesc dd = *desc;
  dd.flags |= GPFLOW_FLAG_FORCE_GENERIC;   // the generic kernel family: the one whose workspace is sized below
  dd.flags &= ~(uint32_t)(GPFLOW_FLAG_FORCE_TC | GPFLOW_FLAG_EULER_PEN);
  size_t need = 0;
  int rc = gpflow_query_workspace(&dd, B < 0 ? 0 : B, GPFLOW_OP_FORWARD, &need);   // validates the descriptor
  if (rc != GPFLOW_OK) return rc;
  if (need > (size_t)GPFLOW_STATIC_SCRATCH_BYTES) return GPFLOW_ERR_STATIC_SCRATCH;
  void* scratch = static_scratch_ptr(&rc);
  if (scratch == nullptr) return rc;
  return gpflow_velocity(&dd, B, params, x, t, v_out, scratch, (size_t)GPFLOW_STATIC_SCRATCH_BYTES, cuda_stream);
}
