"""Host-side mirror of the GP-flow parts of `src/gp_flows/network.py` (class Network) without Lightning.

`gp_flow` (network.py:156-189), `loss_gp_flow` (:70-84), `ot_map` (:191-253) keep their names, arguments and
semantics; `fused_training_step` is the B200 fast path for the GP branch of `training_step` (:104-114): one fused
CUDA call produces the OT cost, the Euler regulariser and the parameter gradients (written to `.grad`), followed --
in data-parallel runs -- by ONE all-reduce of a [P+3] buffer over NCCL.
The frozen normalising-flow map (`self.map`, BNAF / FFJORD / CPFlow modules of the reference) is whatever module the
caller passes in: it runs under no_grad upstream of the hot path and is out of scope here.
"""
from __future__ import annotations

from contextlib import ExitStack
from typing import Optional

import numpy as np
import torch

from . import _lib, data_parallel, ops
import weakref

from .res_net import Compose

_TC_GRAPHS = weakref.WeakKeyDictionary()   # GPFlowNetwork -> captured train steps of the tensor-core path


class GPFlowNetwork(torch.nn.Module):
    def __init__(self, flow, map_module=None, sigma=1.0, process_group=None):
        super().__init__()
        self.map = map_module
        self.flow = flow
        self.sigma = sigma
        self.target_gaussian = True
        self.process_group = process_group
        self._buf = None
        self._flat_opt = None

    # ---- network.py:307-320 (GP branch of configure_optimizers) --------------------------------
    def configure_fused_optimizer(self, lr=2e-3, weight_decay=1e-4, betas=(0.9, 0.999), eps=1e-8, step_size=None,
                                  gamma=0.5):
        """Adam(+L2) + StepLR of the reference on ONE flat parameter buffer (`gpflow_b200.optim.FlatAdam`):
        `fused_training_step` then reads the parameters and writes `.grad` in place, and `opt.step()` is one launch."""
        from .optim import FlatAdam
        self._flat_opt = FlatAdam(self.flow.parameters(), lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                                  step_size=step_size, gamma=gamma)
        return self._flat_opt

    # ---- network.py:156-189 ------------------------------------------------------------------
    def gp_flow(self, x, reverse=False, save_trajectories=False):
        if self.flow is None:
            return (x, torch.zeros_like(x).sum(1)) if not reverse else x
        if isinstance(self.flow, Compose):
            # nb_blocks != 1 (compose.py:5-69): erf once, every block in cube space, clip / erfinv once -- the literal
            # sequence of network.py:174-182 (the blocks' kernels run without their erf / erfinv wrapper).
            eps = ops.eps_of(x.dtype)
            g = torch.erf(x / (np.sqrt(2.0) * self.sigma))
            g = self.flow(g, reverse=reverse, save_trajectories=save_trajectories)
            y = np.sqrt(2) * self.sigma * torch.erfinv(torch.clip(g, -1 + eps, 1 - eps))
        else:
            y = self.flow(x, reverse=reverse, save_trajectories=save_trajectories, _wrap=True, _sigma=self.sigma)
        if reverse:
            return y
        log_abs_det = ((y ** 2).sum(1) - (x ** 2).sum(1)) / (2.0 * self.sigma ** 2)
        return y, log_abs_det

    def _map_points(self, x):
        """`self.map(x)[0]` (network.py:76-78); a map mirror that can skip its log-determinant (FFJORDMap.map_only: no
        trace pass) is asked for the points alone."""
        fn = getattr(self.map, "map_only", None)
        return fn(x) if callable(fn) else self.map(x)[0]

    # ---- network.py:70-84 ---------------------------------------------------------------------
    def loss_gp_flow(self, x, x_map: Optional[torch.Tensor] = None):
        if x_map is None:
            with torch.no_grad():
                x_map = self._map_points(x)
        y, _ = self.gp_flow(x_map)
        return ((x - y) ** 2).sum(-1).mean()

    # ---- fused GP branch of training_step (network.py:104-111) ----------------------------------
    @torch.no_grad()
    def fused_training_step(self, x, x_map: Optional[torch.Tensor] = None, global_batch: Optional[int] = None,
                            with_lagrangian=False):
        """Computes loss_gp = OT cost, reg, and d(loss_gp + reg)/d(flow parameters) into `.grad`.
        Returns (loss_gp, reg) as 0-dim tensors (no host sync).  With a process group the [grads | scalars] buffer is
        all-reduced (SUM) once; `global_batch` defaults to B * world_size."""
        if x_map is None:
            x_map = self._map_points(x)
        flow = self.flow
        world = 1
        pg = self.process_group
        if pg is not None or (torch.distributed.is_available() and torch.distributed.is_initialized()):
            world = torch.distributed.get_world_size(pg)
        B = x.shape[0]
        gb = global_batch if global_batch is not None else B * world
        if isinstance(flow, Compose):
            return self._composed_training_step(x, x_map, gb, world, pg)
        cfg = flow.config(self.sigma)
        euler = flow.euler_coef_penalization if flow._euler_active(x.dtype) else None
        zw = flow.draw_probes(cfg.n_eval, x) if euler is not None else None
        if world > 1:
            zw = data_parallel.broadcast_probes(zw, 0, pg)   # z, w are shared by all samples (res_net.py:330-343)
        P = cfg.n_params
        opt = self._flat_opt
        in_place = opt is not None and opt.n == P and opt.dtype == x.dtype and opt.intact()
        if in_place:            # parameters and .grad are views of the optimiser's flat buffers: no cat, no copies
            flat, buf = opt.flat, opt.gradbuf
        else:
            if self._buf is None or self._buf.numel() != P + _lib.NUM_SCALARS or self._buf.dtype != x.dtype or \
                    self._buf.device != x.device:
                self._buf = torch.empty(P + _lib.NUM_SCALARS, dtype=x.dtype, device=x.device)
            flat, buf = flow.flat_parameters(), self._buf
        res = None
        if in_place and euler is None and not with_lagrangian and not flow.force_generic:
            res = self._graphed_train_step(cfg, flat, x, x_map, 1.0 / gb, buf)
        if res is None:
            res = ops.fused_train_step(cfg, flat, x, x_map, euler_coef=euler, zw=zw, lagrangian=with_lagrangian,
                                       loss_scale=1.0 / gb, out=buf, force_generic=flow.force_generic)
        if world > 1:
            data_parallel.allreduce_train_buffer(res["buf"], pg)
        off = 0
        for p in (() if in_place else flow.parameters()):
            n = p.numel()
            g = res["grads"][off:off + n].view_as(p)
            if p.grad is None:
                p.grad = g.clone()
            else:
                p.grad.copy_(g)
            off += n
        # fresh tensors, as the reference returns (the [grads | scalars] buffer is reused by the next step and mutated in
        # place by the all-reduce; a caller that keeps losses across steps -- Lightning's self.log -- must not alias it)
        sc = res["scalars"].clone()
        flow.regularization_value = sc[_lib.SCALAR_REG]
        if with_lagrangian:
            flow.lagrangian = sc[_lib.SCALAR_LAGR]
        return sc[_lib.SCALAR_OT], sc[_lib.SCALAR_REG]

    def _graphed_train_step(self, cfg, flat, x, x_map, loss_scale, buf):
        """Narrow nets on the tensor-core path enqueue ~2 200 short launches per step (C3 at 2^14 samples: 51.8 ms, of which
        ~4 ms are gaps between launches; at B = 3072 21.1 ms against 16.4 ms).  With stable parameter / gradient buffers
        (the fused optimiser's) the call is captured ONCE per (shape, batch, buffers) into a CUDA graph and replayed; the
        step's inputs are copied into the graph's static input buffers.  Returns None when the step does not qualify
        (GPFLOW_TC_GRAPH=0, other kernel families, wide nets whose launches are long anyway) or the capture failed."""
        import os
        if os.environ.get("GPFLOW_TC_GRAPH", "1") == "0" or x.dtype != torch.float32 or max(cfg.hidden) > 64:
            return None
        if x.shape[0] < 2400 or "gpflow_tc" not in ops.kernel_variant(cfg, x.dtype):
            return None
        if torch.cuda.is_current_stream_capturing():      # the caller is capturing the whole step: plain launches go into ITS graph
            return None
        cache = _TC_GRAPHS.setdefault(self, {})   # outside the module's __dict__: deepcopy / pickle stay clean
        key = (cfg, x.shape[0], x.device, flat.data_ptr(), buf.data_ptr(), float(loss_scale))
        ent = cache.get(key)
        if ent is None:
            try:
                xd, xm = torch.empty_like(x), torch.empty_like(x_map)
                xd.copy_(x)
                xm.copy_(x_map)
                ops.fused_train_step(cfg, flat, xd, xm, loss_scale=loss_scale, out=buf)      # warm-up: kernel attributes, allocator
                graph = torch.cuda.CUDAGraph()
                n0 = int(_lib.load().gpflow_tc_launches_total())
                # thread_local: other threads (an NCCL watchdog, a data loader) are not policed by this capture
                with torch.cuda.graph(graph, capture_error_mode="thread_local"):
                    res = ops.fused_train_step(cfg, flat, xd, xm, loss_scale=loss_scale, out=buf)
                ent = (graph, xd, xm, res, int(_lib.load().gpflow_tc_launches_total()) - n0)
            except Exception as e:  # noqa: BLE001 -- e.g. a capture already in progress: keep the eager CUDA path
                import warnings
                warnings.warn("gpflow_b200: CUDA-graph capture of the train step failed (%r); using eager launches" % (e,))
                ent = False
            if len(cache) >= 8:
                cache.pop(next(iter(cache)))
            cache[key] = ent
        if ent is False:
            return None
        graph, xd, xm, res, n_launch = ent
        xd.copy_(x)
        xm.copy_(x_map)
        graph.replay()
        ops.REPLAYED_TC_LAUNCHES[0] += n_launch
        return res

    def _composed_training_step(self, x, x_map, gb, world, pg):
        """nb_blocks != 1: the blocks are chained through autograd (each block's forward / backward is one fused CUDA op,
        `_FlowFunction`); loss mean over the GLOBAL batch, gradients summed over ranks."""
        params = [q for q in self.flow.parameters()]
        for q in params:
            q.grad = None
        with torch.enable_grad():
            y, _ = self.gp_flow(x_map)
            loss = ((x - y) ** 2).sum(-1).sum() / gb
            reg = self.flow.regularization_coef()
            (loss + reg).backward()
        loss = loss.detach()
        reg = torch.as_tensor(reg, dtype=x.dtype, device=x.device).detach().clone()
        if world > 1:
            flat = torch.cat([q.grad.reshape(-1) for q in params] + [loss.reshape(1)])
            torch.distributed.all_reduce(flat, group=pg)
            off = 0
            for q in params:
                q.grad.copy_(flat[off:off + q.numel()].view_as(q))
                off += q.numel()
            loss = flat[-1].clone()
        return loss, reg

    # ---- network.py:191-253 -------------------------------------------------------------------
    def ot_map(self, x, reverse=False, no_grad_map=False, no_grad_gp_flow=False, return_log_det=False):
        def map_fun(v, reverse):
            return self.map(v) if not reverse else (self.map.backward(v), None)

        def gp_flow_fun(v, reverse):
            return self.gp_flow(v) if not reverse else (self.gp_flow(v, reverse=True), None)

        target_gaussian = (not self.target_gaussian) if reverse else self.target_gaussian
        with ExitStack() as stack:
            if target_gaussian:
                if no_grad_map:
                    stack.enter_context(torch.no_grad())
                map_x, log_map = map_fun(x, reverse)
                stack.close()
                if no_grad_gp_flow:
                    stack.enter_context(torch.no_grad())
                out, log_gp = gp_flow_fun(map_x, reverse)
                stack.close()
            else:
                if no_grad_gp_flow:
                    stack.enter_context(torch.no_grad())
                out, log_gp = gp_flow_fun(x, reverse)
                stack.close()
                if no_grad_map:
                    stack.enter_context(torch.no_grad())
                out, log_map = map_fun(out, reverse)
                stack.close()
        if return_log_det:
            return out, log_map + log_gp
        return out

    def predict_map(self, x, reverse=False):
        return self._map_points(x) if not reverse else self.map.backward(x)

    def get_flow_trajectories(self):
        return self.flow.get_trajectories()
