// gpflow_b200 -- the multi-GPU hook of the C ABI (SURVEY.md 8b "Multi-GPU hook", 8e "Collective"): ONE in-place SUM all-reduce of
// the fused [grads | ot, reg, lagr] buffer on the caller's stream, so that a host that is not PyTorch (or does not want a
// torch.distributed process group) can shard the batch over the GPUs of a box exactly like gpflow_b200/data_parallel.py does.
//
// The reference is single-device (training.py:124-136): there is no reference counterpart to restate; the arithmetic is
// the sum over ranks of per-rank sums already scaled by 1 / B_global (data_parallel.py).
//
// NCCL is bound at RUN time (dlopen of libnccl.so.2, symbols by name): the single-GPU ABI keeps no NCCL dependency, the
// library still links against the CUDA runtime only, and a process that never calls these entry points never loads NCCL.
// Inside a PyTorch process the already-loaded NCCL (torch's own) is found first, so both use the same library.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "../../include/gpflow_b200.h"

namespace {

// the part of nccl.h this file needs (NCCL 2.x ABI: stable enum values and signatures)
typedef struct { char internal[128]; } NcclUniqueId;
typedef void* NcclComm;
enum { kNcclSuccess = 0, kNcclSum = 0, kNcclFloat32 = 7, kNcclFloat64 = 8 };
typedef int (*GetUniqueIdFn)(NcclUniqueId*);
typedef int (*CommInitRankFn)(NcclComm*, int, NcclUniqueId, int);
typedef int (*AllReduceFn)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t);
typedef int (*CommDestroyFn)(NcclComm);

struct NcclApi {
  void* handle = nullptr;
  GetUniqueIdFn get_unique_id = nullptr;
  CommInitRankFn comm_init_rank = nullptr;
  AllReduceFn all_reduce = nullptr;
  CommDestroyFn comm_destroy = nullptr;
  bool ok = false;
};

const NcclApi& nccl_api() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    const char* names[3] = {getenv("GPFLOW_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (int pass = 0; pass < 2 && api.handle == nullptr; ++pass)     // pass 0: a copy that is already loaded (torch's)
      for (const char* n : names) {
        if (n == nullptr || n[0] == 0) continue;
        api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL | (pass == 0 ? RTLD_NOLOAD : 0));
        if (api.handle != nullptr) break;
      }
    if (api.handle == nullptr) return;
    api.get_unique_id = reinterpret_cast<GetUniqueIdFn>(dlsym(api.handle, "ncclGetUniqueId"));
    api.comm_init_rank = reinterpret_cast<CommInitRankFn>(dlsym(api.handle, "ncclCommInitRank"));
    api.all_reduce = reinterpret_cast<AllReduceFn>(dlsym(api.handle, "ncclAllReduce"));
    api.comm_destroy = reinterpret_cast<CommDestroyFn>(dlsym(api.handle, "ncclCommDestroy"));
    api.ok = api.get_unique_id && api.comm_init_rank && api.all_reduce && api.comm_destroy;
  });
  return api;
}

}  // namespace

extern "C" {

int gpflow_comm_unique_id(void* id_out_128_bytes) {
  if (id_out_128_bytes == nullptr) return GPFLOW_ERR_NULL_POINTER;
  const NcclApi& api = nccl_api();
  if (!api.ok) return GPFLOW_ERR_COLLECTIVE;
  NcclUniqueId id;
  if (api.get_unique_id(&id) != kNcclSuccess) return GPFLOW_ERR_COLLECTIVE;
  memcpy(id_out_128_bytes, &id, sizeof(id));
  return GPFLOW_OK;
}

int gpflow_comm_create(const void* id_128_bytes, int n_ranks, int rank, void** comm_out) {
  if (id_128_bytes == nullptr || comm_out == nullptr) return GPFLOW_ERR_NULL_POINTER;
  if (n_ranks < 1 || rank < 0 || rank >= n_ranks) return GPFLOW_ERR_BAD_DESC;
  const NcclApi& api = nccl_api();
  if (!api.ok) return GPFLOW_ERR_COLLECTIVE;
  NcclUniqueId id;
  memcpy(&id, id_128_bytes, sizeof(id));
  NcclComm comm = nullptr;
  if (api.comm_init_rank(&comm, n_ranks, id, rank) != kNcclSuccess) return GPFLOW_ERR_COLLECTIVE;
  *comm_out = comm;
  return GPFLOW_OK;
}

int gpflow_allreduce_grads(void* buf, int64_t count, int dtype, void* comm, void* cuda_stream) {
  if (buf == nullptr || comm == nullptr) return GPFLOW_ERR_NULL_POINTER;
  if (count < 0 || (dtype != GPFLOW_F32 && dtype != GPFLOW_F64)) return GPFLOW_ERR_BAD_DESC;
  if (count == 0) return GPFLOW_OK;
  const NcclApi& api = nccl_api();
  if (!api.ok) return GPFLOW_ERR_COLLECTIVE;
  const int rc = api.all_reduce(buf, buf, (size_t)count, dtype == GPFLOW_F64 ? kNcclFloat64 : kNcclFloat32, kNcclSum,
                                reinterpret_cast<NcclComm>(comm), reinterpret_cast<cudaStream_t>(cuda_stream));
  return rc == kNcclSuccess ? GPFLOW_OK : GPFLOW_ERR_COLLECTIVE;
}

int gpflow_comm_destroy(void* comm) {
  if (comm == nullptr) return GPFLOW_OK;
  const NcclApi& api = nccl_api();
  if (!api.ok) return GPFLOW_ERR_COLLECTIVE;
  return api.comm_destroy(reinterpret_cast<NcclComm>(comm)) == kNcclSuccess ? GPFLOW_OK : GPFLOW_ERR_COLLECTIVE;
}

}  // extern "C"
