"""world_size-2 gloo tests (CPU) of the data-parallel host logic: batch shards, loss_scale = 1/B_global,
ONE all-reduce(SUM) of the fused [grads | ot, reg, lagr] buffer, shared Euler probes.  The per-rank buffers are
produced by the CPU oracle here (the checker standing in for the kernel, which needs a GPU); what is under test is
the sharding / reduction contract the CUDA path relies on (the GPU counterpart is
tests/test_sf2_parity.py::test_sf2_buffer_is_additive_over_batch_shards)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import gpflow_oracle as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_buffer(params, x_data, x_map, spec, B_global, euler, zw):
    """[grads | ot, reg, 0] of one shard with sums scaled by 1/B_global (what gpflow_train_step returns)."""
    P = sum(p.numel() for p in params)
    buf = torch.zeros(P + 3, dtype=torch.float64)
    n = x_map.shape[0]
    if n == 0:
        return buf
    r = O.train_step(params, x_data, x_map, spec, euler_coef=euler, zw=zw)
    w = n / B_global                      # oracle returns shard MEANS; the kernel returns SUMS / B_global
    buf[:P] = r["grads"] * w
    buf[P] = r["ot"] * w
    buf[P + 1] = r["reg"] * w
    return buf


def _worker(rank, world, port, B, euler, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from gpflow_b200 import data_parallel as DP
        torch.set_num_threads(1)
        spec = O.FlowSpec(d=2, nb=1, hidden=[15, 15], n_steps=4)
        params = O.init_params(spec, 0.3, 0, torch.float64)
        g = torch.Generator().manual_seed(7)
        x_map = torch.randn(B, 2, generator=g, dtype=torch.float64)
        x_data = 1.2 * x_map.flip(-1) + 0.3
        zw = None
        if euler is not None:
            # rank 0 draws the probes (z first, then w, per step: res_net.py:330-333); the others receive them
            zw = torch.randn(spec.n_steps, 2, 2, generator=torch.Generator().manual_seed(100 + rank), dtype=torch.float64)
            zw = DP.broadcast_probes(zw, 0)
        lo, hi = DP.shard_bounds(B, rank, world)
        buf = _rank_buffer(params, x_data[lo:hi], x_map[lo:hi], spec, B, euler, zw)
        DP.allreduce_train_buffer(buf)
        if rank == 0:
            zw0 = None
            if euler is not None:
                zw0 = torch.randn(spec.n_steps, 2, 2, generator=torch.Generator().manual_seed(100), dtype=torch.float64)
            full = _rank_buffer(params, x_data, x_map, spec, B, euler, zw0)
            torch.save({"reduced": buf, "full": full}, out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("B,euler", [(64, None), (37, None), (1, None), (48, 5e-3)])
def test_two_rank_allreduce_reproduces_global_batch(tmp_path, B, euler):
    out = str(tmp_path / "r.pt")
    mp.spawn(_worker, args=(2, _free_port(), B, euler, out), nprocs=2, join=True)
    r = torch.load(out)
    err = (r["reduced"] - r["full"]).norm() / r["full"].norm()
    # the Euler penalty is a finite difference with delta = 2 sqrt(eps64) = 3e-8 (res_net.py:346): its own rounding
    # noise (~1e-8 relative) differs between the sharded and the full evaluation
    assert err < (1e-12 if euler is None else 1e-7), float(err)


def test_shard_bounds_cover_the_batch():
    from gpflow_b200.data_parallel import shard_bounds
    for n in (0, 1, 7, 8, 1 << 24, (1 << 24) + 5):
        for world in (1, 2, 3, 4, 8):
            cuts = [shard_bounds(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(cuts[:-1], cuts[1:]))
            sizes = [hi - lo for lo, hi in cuts]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_bounds(4, 2, 2)
