"""Host -> device input staging for the GP-flow train step (the step either side of the path).

The reference feeds the step from a CPU `Dataset` through a DataLoader (`src/data_module.py:146-186`): every step
starts with a blocking host->device copy of `(x, x_map)`.  `HostBatchStream` keeps that interface (iterate, get device
tensors) but double-buffers: batch k+1 is copied from pinned host memory on a side CUDA stream while the kernels of
batch k run, and an event hands the buffers over, so the copy (268 MB per 2^24-sample batch) leaves the critical path.
Plumbing only (torch streams / events / pinned memory); no arithmetic of the path happens here.
"""
from __future__ import annotations

from typing import Iterable, Iterator, Sequence, Tuple

import torch


class HostBatchStream:
    def __init__(self, host_batches: Iterable[Sequence[torch.Tensor]], device, depth: int = 2):
        """`host_batches` yields tuples of CPU tensors (pinned for truly asynchronous copies); `depth` device buffer
        sets are rotated (2 = classic double buffering)."""
        if depth < 2:
            raise ValueError("depth must be >= 2")
        self.src = host_batches
        self.device = torch.device(device)
        self.depth = depth
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.bytes_copied = 0

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, ...]]:
        it = iter(self.src)
        slots = [None] * self.depth      # per slot: (device tensors, copy-done event, consumer-done event)
        main = torch.cuda.current_stream(self.device)

        def stage(i, host):
            bufs, _, consumed = slots[i] if slots[i] is not None else (None, None, None)
            if bufs is None or any(b.shape != h.shape or b.dtype != h.dtype for b, h in zip(bufs, host)):
                bufs = tuple(torch.empty(h.shape, dtype=h.dtype, device=self.device) for h in host)
                consumed = None
            with torch.cuda.stream(self.copy_stream):
                if consumed is not None:
                    self.copy_stream.wait_event(consumed)        # the kernels that read this slot have finished
                for b, h in zip(bufs, host):
                    b.copy_(h, non_blocking=True)
                    self.bytes_copied += h.numel() * h.element_size()
                done = torch.cuda.Event()
                done.record(self.copy_stream)
            slots[i] = (bufs, done, None)

        k = 0
        try:
            stage(0, next(it))
        except StopIteration:
            return
        while True:
            cur = k % self.depth
            nxt = None
            try:
                nxt = next(it)
            except StopIteration:
                pass
            if nxt is not None:
                stage((k + 1) % self.depth, nxt)                  # overlaps with the step the caller runs on `cur`
            bufs, done, _ = slots[cur]
            main.wait_event(done)
            yield bufs
            consumed = torch.cuda.Event()
            consumed.record(main)                                 # everything the caller enqueued on this batch
            slots[cur] = (bufs, done, consumed)
            if nxt is None:
                return
            k += 1
