"""Data-parallel plumbing of the GP-flow train step: one process per GPU, batch shards, ONE all-reduce.

The path shards naturally: samples are independent through erf -> RK4 -> erfinv, the only coupling is the batch mean
of the OT cost / Euler penalty / Lagrangian (network.py:82, res_net.py:293,358) and the shared parameters.  Each rank
therefore runs the fused kernel on its contiguous shard with `loss_scale = 1 / B_global`; the kernel's output is one
contiguous buffer `[grads (P) | ot, reg, lagr]` of *sums already scaled by 1/B_global*, so a single
`all_reduce(SUM)` of P + 3 elements (335 floats = 1.3 KB for the 2-D flow) yields the global-batch loss and gradient
on every rank.  The Euler probes z, w are shared by all samples of a step (res_net.py:330-343): every rank must
draw them from the same generator state (see `broadcast_probes`).

Nothing here touches the arithmetic of the path: it is host logic over `torch.distributed` (NCCL on the GPUs,
gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of a batch of n samples for `rank` of `world`; the first n % world ranks get one
    extra sample, so ragged batches are covered (empty shards are legal when n < world)."""
    if world < 1 or not (0 <= rank < world) or n < 0:
        raise ValueError("bad shard request n=%d rank=%d world=%d" % (n, rank, world))
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def world_size(group=None) -> int:
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(group)
    return 1


def allreduce_train_buffer(buf: torch.Tensor, group=None) -> torch.Tensor:
    """In-place SUM all-reduce of the fused [grads | ot, reg, lagr] buffer.  No-op in single-process runs."""
    if world_size(group) > 1:
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    return buf


def broadcast_probes(zw: Optional[torch.Tensor], src: int = 0, group=None) -> Optional[torch.Tensor]:
    """Make the Euler probes identical on every rank (rank `src` drew them with the reference's RNG order)."""
    if zw is not None and world_size(group) > 1:
        dist.broadcast(zw, src=src, group=group)
    return zw


def split_buffer(buf: torch.Tensor, n_params: int):
    """(grads [P], ot, reg, lagr) views of the fused buffer."""
    return buf[:n_params], buf[n_params], buf[n_params + 1], buf[n_params + 2]


class NativeAllReduce:
    """The C ABI's own all-reduce (`gpflow_allreduce_grads`, include/gpflow_b200.h "Multi-GPU hook"): an NCCL communicator
    created by the library itself (NCCL bound at run time), bootstrapped over an existing torch.distributed group -- rank 0's
    128-byte id is broadcast through it -- or, for a single process, with n_ranks = 1.  `allreduce_(buf)` enqueues the
    in-place SUM on the current CUDA stream, like `allreduce_train_buffer`.  A host without torch.distributed calls the four
    C entry points directly and moves the id by its own means."""

    def __init__(self, device, group=None):
        import ctypes
        from . import _lib
        self._lib = _lib.load()
        self.device = torch.device(device)
        world = world_size(group)
        rank = dist.get_rank(group) if world > 1 else 0
        ident = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            _check(self._lib.gpflow_comm_unique_id(ctypes.c_void_p(ident.data_ptr())), "gpflow_comm_unique_id")
        if world > 1:
            backend = dist.get_backend(group)
            box = ident.to(self.device) if backend == "nccl" else ident
            dist.broadcast(box, src=0, group=group)
            ident = box.cpu()
        comm = ctypes.c_void_p(0)
        with torch.cuda.device(self.device):
            _check(self._lib.gpflow_comm_create(ctypes.c_void_p(ident.data_ptr()), world, rank, ctypes.byref(comm)),
                   "gpflow_comm_create")
        self._comm, self.world, self.rank = comm, world, rank

    def allreduce_(self, buf: torch.Tensor) -> torch.Tensor:
        import ctypes
        from . import _lib
        if not buf.is_cuda or not buf.is_contiguous() or buf.dtype not in (torch.float32, torch.float64):
            raise ValueError("NativeAllReduce: contiguous fp32 / fp64 CUDA tensor expected")
        with torch.cuda.device(buf.device):
            stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
            _check(self._lib.gpflow_allreduce_grads(ctypes.c_void_p(buf.data_ptr()), buf.numel(),
                                                    _lib.F64 if buf.dtype == torch.float64 else _lib.F32, self._comm, stream),
                   "gpflow_allreduce_grads")
        return buf

    def close(self):
        if self._comm is not None and self._comm.value:
            self._lib.gpflow_comm_destroy(self._comm)
        self._comm = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001 -- interpreter shutdown
            pass


def _check(rc: int, what: str):
    if rc != 0:
        from . import _lib
        raise RuntimeError("%s failed: %s" % (what, _lib.load().gpflow_error_string(rc).decode()))
