#!/usr/bin/env python
"""bench.py -- GP-flow train-step throughput (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W]            # our CUDA path
    python bench.py --impl reference [--gpus N] [--steps K] ...    # CPU reference arm (oracle port, host cores)

Workload (BASELINE.json configs[3], SURVEY.md 8d "C4"): synthetic 2-D Gaussian-noise train_gp step, global batch
2^24 samples sharded over the ranks (strong scaling), d=2, nb_div_blocks=1, velocity MLP [15,15], 15 RK4 steps,
fp32.  One step = fused forward + OT cost + backward (one CUDA launch + deterministic finalize) -> ONE all-reduce of
the [grads | scalars] buffer (N > 1) -> fused Adam update (gpflow_adam_step) of the flat parameter buffer.
Prints ONE JSON line (see the repo's task contract for the keys).  The default run appends `rows` to that line: the other
BASELINE.json configurations -- C5 (d=64, nb=63, H=[256,256]) at 2^14 samples PER GPU and C3 (d=10, nb=9, H=[50,50,50]) on
the chunked tcgen05 path (weak scaling), C4 with the Euler penalization -- each with its own roofline / clocks / e2e, and
`peaks`: the roofline denominators measured once, before any warm-up (FFMA burn + SM clock, torch.matmul TF32 8192^3).
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (d, nb, hidden, n_steps, log2 global batch, weight std)
    "c4_2d_2p24": dict(d=2, nb=1, hidden=(15, 15), n_steps=15, log2_batch=24, w_std=0.3,
                       desc="synthetic train_gp_on_gaussian_noise 2D, 2^24-sample global batch"),
    "c3_d10_nb9": dict(d=10, nb=9, hidden=(50, 50, 50), n_steps=15, log2_batch=14, w_std=0.05,
                       desc="dsprites-latent shape d=10 nb=9 H=[50,50,50] (chunked tcgen05 path, widths padded to 64)"),
    "c5_d64_nb63": dict(d=64, nb=63, hidden=(256, 256), n_steps=15, log2_batch=14, w_std=0.02,
                        desc="synthetic d=64 nb=63 H=[256,256], 2^14 samples per GPU (chunked tcgen05 path)"),
}


def synth_batch(B, d, seed, device):
    """x_map ~ N(0,I); x_data = deterministic affine image (stand-in for map.backward)."""
    import torch
    g = torch.Generator(device=device).manual_seed(seed)
    x_map = torch.randn(B, d, generator=g, device=device, dtype=torch.float32)
    th = math.pi / 3
    A = torch.eye(d, device=device)
    A[0, 0], A[0, 1], A[1, 0], A[1, 1] = math.cos(th), -math.sin(th), math.sin(th), math.cos(th)
    scale = torch.ones(d, device=device)
    scale[0], scale[1] = 1.5, 0.7
    shift = torch.zeros(d, device=device)
    shift[0], shift[1] = 0.5, -0.25
    x_data = (x_map @ A) * scale + shift
    return x_data.contiguous(), x_map.contiguous()


class ClockSampler:
    """Clocks / throttle reasons DURING the timed region (B200_PROFILING.md).
    * throttle reasons + clocks.sm through NVML (the counters `nvidia-smi --query-gpu=clocks.sm,clocks_event_reasons.*`
      prints) from a thread of this process, every 100 ms;
    * the SM clock under load from an on-device probe (gpflow_clock_probe: clock64 ticks over globaltimer ns) enqueued
      after every timed step -- no driver round trip, one sample per step."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index):
        self.index, self.sm, self.bits, self.th, self.stop_flag, self.smax, self.err = index, [], 0, None, False, None, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[self.index]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else self.index
            h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.smax = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))

            def loop():
                while not self.stop_flag:
                    try:
                        self.sm.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                        self.bits |= int(pynvml.nvmlDeviceGetCurrentClocksEventReasons(h))
                    except Exception as e:  # noqa: BLE001
                        self.err = repr(e)
                    time.sleep(0.1)

            self.th = threading.Thread(target=loop, daemon=True)
            self.th.start()
        except Exception as e:  # noqa: BLE001
            self.err = repr(e)

    def stop(self):
        self.stop_flag = True
        if self.th is not None:
            self.th.join(timeout=1.0)
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.smax, "reasons": ["nvml unavailable: %s" % self.err]}
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2], "sm_min_mhz": sm[0], "sm_max_mhz": self.smax,
                "reasons": sorted(v for k, v in self.REASONS.items() if self.bits & k), "nvml_samples": len(sm)}


def cpu_oracle_rate(wl, sample, steps, warmup, threads):
    """The reference's algorithm on the host cores: oracle port (torch CPU, autograd), loss_gp_flow + backward."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import gpflow_oracle as O
    torch.set_num_threads(threads)
    spec = O.FlowSpec(d=wl["d"], nb=wl["nb"], hidden=list(wl["hidden"]), n_steps=wl["n_steps"])
    params = O.init_params(spec, wl["w_std"], 0, torch.float32)
    x_data, x_map = synth_batch(sample, wl["d"], 1234, "cpu")
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        O.train_step(params, x_data, x_map, spec)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    return sample * len(times) / sum(times), sum(times) / len(times)


def run_reference(args, wl, rank, world):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    sample = 1 << args.ref_log2_sample
    rate, sec = cpu_oracle_rate(wl, sample, args.steps, args.warmup, threads)
    line = {
        "impl": "reference", "metric": "gp_flow_train_step_samples_per_sec", "value": rate, "unit": "samples/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload + ": " + wl["desc"], "sample_per_step": sample},
        "cpu_baseline": {"value": rate, "unit": "samples/s", "cores": threads, "kind": "port",
                         "sample": "%d samples/step of the same synthetic batch; reference is pure Python and cannot "
                                   "travel to the GPU box, so the oracle port (torch CPU autograd) is timed" % sample},
        "e2e": {"value": rate, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def run_ffjord_dataset(args, rank, world, local_rank):
    """Extra row: the Gaussian-noise data set of the latent configs, `map.backward(gauss)` through
    the frozen pretrained dsprites FFJORD (src/data_module.py:161-186), one gpflow_ffjord_integrate launch per step.
    A step = one data-set build of 2^log2_batch points (default 2^18 ~ the reference's 240 000); ranks build independent
    shards (no collective).  e2e adds the device->host copy of the data set (the reference keeps it in a CPU Dataset)."""
    json_fd = os.dup(1)
    os.dup2(2, 1)
    import numpy as np
    import torch
    import torch.distributed as dist
    import gpflow_b200 as G
    from gpflow_b200 import ops

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    z = np.load(os.path.join(ROOT, "tests", "golden", "ffjord_dsprites_f32.npz"))
    d, hidden, nblocks, n_steps = int(z["d"]), [int(h) for h in z["hidden"]], int(z["n_blocks"]), 64
    t_end = [float(t) for t in z["t_end"]]
    m = G.FFJORDMap(d, hidden, nblocks, torch.from_numpy(z["params"]), t_end, n_steps).to(dev)
    B_global = 1 << (args.log2_batch or 18)
    B = B_global // world
    n_buf = 8                                   # 8 x 10 MB of distinct Gaussian points; the kernel is compute-bound (80 B/sample)
    gauss = [torch.randn(B, d, device=dev, generator=torch.Generator(device=dev).manual_seed(1234 + rank + 1000 * i))
             for i in range(n_buf)]
    for i in range(args.warmup):
        m.backward(gauss[i % n_buf])
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(args.steps):
        x = m.backward(gauss[i % n_buf])
    ev1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t = torch.tensor([ev0.elapsed_time(ev1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0]) / args.steps
    clk = clocks.stop() if rank == 0 else None
    # e2e: pinned HOST Gaussian points in, data set back on the host (the reference's Dataset lives on the CPU)
    hin = [g.cpu().pin_memory() for g in gauss[:2]]
    hout = torch.empty(B, d).pin_memory()
    def e2e_run(n):
        for i in range(n):
            hout.copy_(m.backward(hin[i % 2].to(dev, non_blocking=True)), non_blocking=True)
        torch.cuda.synchronize()
    e2e_run(1)
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); e2e_run(args.steps); e1.record(); torch.cuda.synchronize()
    te = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    n = [d] + hidden + [d]
    flops = 2.0 * sum(n[i] * n[i + 1] for i in range(len(n) - 1)) * 4 * n_steps * nblocks      # per sample, value only
    peak_tf = ops.fp32_fma_peak_tflops()
    achieved = flops * B / (ms * 1e-3) / 1e12
    cpu_baseline = None
    if not args.no_cpu_baseline and world == 1:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))      # CPU baseline leg only
        import ffjord_oracle as FO
        threads = os.cpu_count() or 1
        torch.set_num_threads(threads)
        spec = FO.FfjordSpec(d=d, hidden=hidden, n_blocks=nblocks, n_steps=n_steps, t_end=t_end)
        xs = torch.randn(1 << 14, d)
        with torch.no_grad():
            FO.integrate(torch.from_numpy(z["params"]), xs[:1024], spec, reverse=True)
            t0 = time.perf_counter()
            FO.integrate(torch.from_numpy(z["params"]), xs, spec, reverse=True)
            sec = time.perf_counter() - t0
        cpu_baseline = {"value": xs.shape[0] / sec, "unit": "samples/s", "cores": threads, "kind": "port",
                        "sample": "%d samples, same map and step count (oracle port, torch CPU, all host threads)" % xs.shape[0]}
    line = {
        "metric": "ffjord_gaussian_noise_dataset_samples_per_sec", "value": B_global / (ms * 1e-3), "unit": "samples/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "ffjord_dataset: map.backward(gauss) through pretrained dsprites FFJORD (3 CNF blocks "
                               "10->64->64->10), %d RK4 steps per block" % n_steps, "global_batch": B_global,
                   "per_gpu_batch": B, "parallelism": "dp%d (independent shards, no collective)" % world,
                   "l2": "compute-bound (80 B/sample); rotating through %d distinct input buffers" % n_buf},
        "clocks": clk,
        "e2e": {"value": B_global / (float(te[0]) / args.steps * 1e-3), "unit": "samples/s",
                "h2d_bytes_per_step": B * d * 4 * world, "d2h_bytes_per_step": B * d * 4 * world,
                "api": "pinned host gauss -> FFJORDMap.backward -> pinned host data set"},
        "gpu_launches": args.steps,
        "roofline": {"bound": "fp32_fma", "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s",
                     "frac": achieved / peak_tf if peak_tf else None, "traffic": None,
                     "kernel": "gpflow_ffjord_kernel<f32>", "kernel_ms": ms, "algorithmic_flops_per_sample": flops,
                     "samples_per_launch": B,
                     "peak_source": "FFMA burn micro-kernel measured live in this run (gpflow_fp32_fma_burn)"},
        "cpu_baseline": cpu_baseline,
    }
    sys.stdout.flush()
    os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c4_2d_2p24", choices=sorted(WORKLOADS) + ["ffjord_dataset"])
    ap.add_argument("--log2-batch", type=int, default=None, help="override the global batch (log2)")
    ap.add_argument("--ref-log2-sample", type=int, default=15)
    ap.add_argument("--cpu-baseline-sample", type=int, default=1 << 17)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--dbg", default="", help=argparse.SUPPRESS)   # developer toggles: nbuf1,noopt,noclk
    ap.add_argument("--euler", type=float, default=None, help="Euler penalization coefficient (extra row)")
    ap.add_argument("--rows", default="all", help="extra rows of the default run: all | none | comma list of " +
                    "c5_d64_nb63,c3_d10_nb9,c4_2d_euler_5e-4")
    args = ap.parse_args()
    wl = dict(WORKLOADS.get(args.workload, {}))
    if args.log2_batch is not None:
        wl["log2_batch"] = args.log2_batch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus > 1 and world == 1 and "RANK" not in os.environ:
        # convenience: re-launch under torchrun exactly as the driver does
        port = 29500 + os.getpid() % 1000
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", str(port)] + sys.argv
        sys.exit(subprocess.call(cmd))

    if args.workload == "ffjord_dataset":
        if args.impl == "reference":
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": "ffjord_dataset is an extra row; its CPU leg is "
                                  "the cpu_baseline object of the normal run"}))
            return
        run_ffjord_dataset(args, rank, world, local_rank)
        return
    if args.impl == "reference":
        run_reference(args, wl, rank, world)
        return

    # NCCL / torchrun may print banners on stdout: route fd 1 to stderr for the run, keep the real one for the JSON line
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    peaks = measure_peaks(dev)            # BEFORE any warm-up: cool chip, max clock (VERDICT r1 "what's weak" 10)
    head = measure_workload(args, args.workload, wl, args.euler, rank, world, local_rank, dev, args.steps, args.warmup,
                            peaks, weak=False, with_cpu_baseline=not args.no_cpu_baseline)
    rows = []
    if args.rows != "none" and args.workload == "c4_2d_2p24" and args.euler is None and args.log2_batch is None:
        for spec in ROWS:
            if args.rows != "all" and spec["name"] not in args.rows.split(","):
                continue
            rwl = dict(WORKLOADS[spec["workload"]])
            rwl["log2_batch"] = spec["log2_batch"]
            r = measure_workload(args, spec["workload"], rwl, spec.get("euler"), rank, world, local_rank, dev, spec["steps"],
                                 spec["warmup"], peaks, weak=spec["weak"], with_cpu_baseline=False)
            if rank == 0:
                r["name"] = spec["name"]
                rows.append(r)
    if rank == 0:
        head["rows"] = rows
        head["peaks"] = peaks
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(head) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


# extra rows of the default run (VERDICT r1 "next round" 1b): the other BASELINE.json configs, each with its own
# roofline / clocks / e2e, so that the driver's BENCH / SCALE files carry them at N = 1, 2, 4, 8
ROWS = [
    # C5 at 2^14 samples PER GPU (SURVEY.md 8d "C5 inputs"): weak scaling
    dict(name="c5_d64_nb63", workload="c5_d64_nb63", log2_batch=14, weak=True, steps=3, warmup=3),
    dict(name="c3_d10_nb9", workload="c3_d10_nb9", log2_batch=14, weak=True, steps=5, warmup=3),
    # C2's penalised step on the C4 data: 2^22 global
    dict(name="c4_2d_euler_5e-4", workload="c4_2d_2p24", log2_batch=22, weak=False, euler=5e-4, steps=3, warmup=3),
]

NOMINAL_FP32_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12      # 74.5: 148 SMs x 128 FP32 lanes x 2 (FMA) x 1.965 GHz


def _sm_mhz_now(dev):
    """SM clock from the on-device probe (clock64 ticks over globaltimer ns), right now on the current stream."""
    import ctypes
    import torch
    from gpflow_b200 import _lib
    out = torch.zeros(2, dtype=torch.int64, device=dev)
    _lib.load().gpflow_clock_probe(ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    o = out.cpu().double()
    return round(float(1e3 * o[0] / max(float(o[1]), 1.0)), 1)


def measure_peaks(dev):
    """Roofline denominators, measured ONCE before any warm-up.
    fp32: FFMA-saturating micro-kernel (gpflow_fp32_fma_burn) with the SM clock seen right after it; nominal 74.5 beside it.
    tf32: dense TF32 GEMM rate as SURVEY.md 8(d) prescribes -- torch.matmul (cuBLAS) TF32 8192^3, best of 10 (burst) and
    back to back for ~2 s (sustained, what a kernel inside a second-long step sees under the power cap).  A 3xTF32
    contraction is judged against a THIRD of it (fp32-equivalent)."""
    import torch
    from gpflow_b200 import ops
    fma = ops.fp32_fma_peak_tflops()
    fma_mhz = _sm_mhz_now(dev)
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    n = 8192
    a = torch.randn(n, n, device=dev)
    b = torch.randn(n, n, device=dev)
    c = torch.empty(n, n, device=dev)
    fl = 2.0 * n ** 3
    torch.matmul(a, b, out=c)
    torch.cuda.synchronize()
    best = 0.0
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); torch.matmul(a, b, out=c); e1.record(); e1.synchronize()
        best = max(best, fl / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    reps = max(8, int(2.0 * best * 1e12 / fl))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        torch.matmul(a, b, out=c)
    e1.record(); e1.synchronize()
    sustained = reps * fl / (e0.elapsed_time(e1) * 1e-3) / 1e12
    tf32_mhz = _sm_mhz_now(dev)
    torch.backends.cuda.matmul.allow_tf32 = old
    del a, b, c
    torch.cuda.empty_cache()
    measured = {}
    try:
        measured = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    return {"fp32_fma_tflops": fma, "fp32_fma_sm_mhz": fma_mhz, "fp32_nominal_tflops": NOMINAL_FP32_TFLOPS,
            "tf32_dense_tflops_burst": best, "tf32_dense_tflops_sustained": sustained, "tf32_sm_mhz_after_sustained": tf32_mhz,
            "how": "before warm-up: gpflow_fp32_fma_burn best of 5; torch.matmul TF32 8192^3 best of 10 (burst) and "
                   "%d back to back (sustained)" % reps,
            "hbm_gbs_measured": measured.get("hbm_gbs"), "bf16_tflops_measured": measured.get("bf16_tflops")}


def measure_workload(args, name, wl, euler, rank, world, local_rank, dev, steps, warmup, peaks, weak, with_cpu_baseline):
    """One row: W warm-up + K timed steps of the fused train step (+ all-reduce + fused Adam) on resident inputs, the same
    through the public API with HOST buffers (e2e), roofline of the dominant kernel family.  weak: per-GPU batch fixed
    at 2^log2_batch (global = world x that); else the global batch is fixed."""
    import ctypes
    import torch
    import torch.distributed as dist
    import gpflow_b200 as G
    from gpflow_b200 import ops, _lib
    from gpflow_b200.roofline import algorithmic_flops_per_sample, executed_flops_per_sample_tc

    B = (1 << wl["log2_batch"]) if weak else (1 << wl["log2_batch"]) // world
    B_global = B * world
    d = wl["d"]
    flow = G.get_gp_flow(d, {"nb_neurons": list(wl["hidden"]), "nb_div_blocks": wl["nb"]}, wl["n_steps"],
                         euler_dict={"use_euler": euler is not None, "case": "penalization",
                                     "coef_penalization": euler}).to(dev)
    torch.manual_seed(0)
    with torch.no_grad():
        for p in flow.parameters():
            p.normal_(0.0, wl["w_std"])           # N(0, w_std^2) weights AND biases, seed 0
    net = G.GPFlowNetwork(flow, map_module=None)
    # network.py:307-320: Adam(lr, betas=(0.9, 0.999), weight_decay=1e-4) on the flat parameter buffer, one launch per step
    opt = net.configure_fused_optimizer(lr=2e-3, weight_decay=1e-4)
    cfg = flow.config()
    lib = _lib.load()

    # inputs resident in HBM; rotate through enough distinct batches to exceed the 126 MB L2
    bytes_per_batch = 2 * B * d * 4
    n_buf = max(1, min(8, math.ceil(300e6 / bytes_per_batch)))
    if "nbuf1" in args.dbg:
        n_buf = 1
    bufs = [synth_batch(B, d, 1234 + rank + 1000 * i, dev) for i in range(n_buf)]

    for i in range(warmup):
        x_data, x_map = bufs[i % n_buf]
        net.fused_training_step(x_data, x_map, global_batch=B_global)
        opt.step()
    torch.cuda.synchronize()
    per_rank = None
    if world > 1 and euler is None:
        # diagnostic, outside the timed region: every rank's kernel time WITHOUT the all-reduce, all GPUs of the box busy at
        # the same time -- the synchronous step runs at the pace of the slowest of these
        dist.barrier()
        flat0 = flow.flat_parameters().detach().clone()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ops.fused_train_step(cfg, flat0, *bufs[0], loss_scale=1.0 / B_global)
        s0.record()
        for i in range(3):
            ops.fused_train_step(cfg, flat0, *bufs[i % n_buf], loss_scale=1.0 / B_global)
        s1.record()
        torch.cuda.synchronize()
        mine = torch.tensor([s0.elapsed_time(s1) / 3.0], device=dev, dtype=torch.float64)
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = {"solo_kernel_ms": [round(float(v), 3) for v in allr],
                    "what": "3 train-step launches per rank without the all-reduce, all ranks at once, before the timed region"}
    if world > 1:
        dist.barrier()
    clocks = ClockSampler(local_rank)
    if rank == 0 and "noclk" not in args.dbg:
        clocks.start()   # no sleep here: a delay on one rank after the barrier would be charged to every rank
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    probe = torch.zeros(steps, 2, dtype=torch.int64, device=dev)
    stream_ptr = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    tc_l0 = ops.tc_launches_total()      # eager launches + launches inside replayed CUDA graphs
    ev0.record()
    for i in range(steps):
        x_data, x_map = bufs[i % n_buf]
        # the fused train step (its kernels + finalize [+ all-reduce]) bracketed by its own events: roofline numerator
        kev[i][0].record()
        loss, reg = net.fused_training_step(x_data, x_map, global_batch=B_global)
        kev[i][1].record()
        if "noopt" not in args.dbg:
            opt.step()
        lib.gpflow_clock_probe(ctypes.c_void_p(probe[i].data_ptr()), stream_ptr)   # ~20 us, SM clock right after the step
    ev1.record()
    torch.cuda.synchronize()
    tc_launches = ops.tc_launches_total() - tc_l0
    if world > 1:
        dist.barrier()
    ms_total = ev0.elapsed_time(ev1)
    ms_kernel = sum(a.elapsed_time(b) for a, b in kev) / steps
    t = torch.tensor([ms_total, ms_kernel], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_kernel = float(t[0]), float(t[1])
    clk = clocks.stop() if rank == 0 else None
    if rank == 0:
        pr = probe.cpu().double()
        mhz = sorted((1e3 * pr[:, 0] / pr[:, 1].clamp(min=1)).tolist())
        clk["nvml_sm_mhz"] = clk.get("sm_mhz")
        clk["sm_mhz"] = round(mhz[len(mhz) // 2], 1)          # median of the on-device probes, one per timed step
        clk["sm_min_mhz"] = round(mhz[0], 1)
        clk["samples"] = len(mhz)
    ms_per_step = ms_total / steps
    value = B_global / (ms_per_step * 1e-3)

    # ---- e2e: public API with HOST buffers; H2D of the step's inputs and D2H of the loss inside the timed region
    e2e = None
    if not args.no_e2e:
        hx = [(a.cpu().pin_memory(), b.cpu().pin_memory()) for a, b in bufs[:min(n_buf, 2)]]

        def e2e_run(n):
            """n steps through the public API: pinned HOST batches -> gpflow_b200.HostBatchStream (H2D of step k+1 on a
            side stream while step k runs) -> GPFlowNetwork.fused_training_step -> Adam -> loss read back to the host"""
            stream = G.HostBatchStream((hx[i % len(hx)] for i in range(n)), dev)
            tot = 0.0
            for dx, dm in stream:
                loss_, reg_ = net.fused_training_step(dx, dm, global_batch=B_global)
                opt.step()
                tot += float(loss_) + float(reg_)      # device->host read of the step's result
            return stream.bytes_copied

        e2e_run(2)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        h2d = e2e_run(steps)
        e1.record()
        torch.cuda.synchronize()
        te = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e = {"value": B_global / (float(te[0]) / steps * 1e-3), "unit": "samples/s",
               "h2d_bytes_per_step": h2d // steps * world, "d2h_bytes_per_step": 8 * world,
               "api": "HostBatchStream -> GPFlowNetwork.fused_training_step -> FlatAdam.step -> float(loss)"}
    if rank != 0:
        return None

    # ---- roofline of the dominant kernel family
    kernel = G.kernel_variant(cfg, euler_coef=euler)
    uses_tc = tc_launches > 0
    flops_alg = algorithmic_flops_per_sample(d, wl["nb"], wl["hidden"], wl["n_steps"], train=True, euler=euler is not None)
    achieved = flops_alg * B / (ms_kernel * 1e-3) / 1e12
    if uses_tc:
        # the wide shapes are judged against the tensor pipe at the precision actually used: 3xTF32 =>
        # a third of the dense TF32 rate.  The step is seconds of back-to-back GEMMs => the SUSTAINED figure.
        flops_exe = executed_flops_per_sample_tc(d, wl["nb"], wl["hidden"], wl["n_steps"], train=True)
        achieved_exe = flops_exe * B / (ms_kernel * 1e-3) / 1e12
        peak = peaks["tf32_dense_tflops_sustained"] / 3.0
        desc = ops.make_desc(cfg, torch.float32, wrap=True)
        roofline = {
            "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
            "achieved_executed": achieved_exe, "frac_executed": achieved_exe / peak,
            "frac_of_burst": achieved / (peaks["tf32_dense_tflops_burst"] / 3.0),
            "traffic": None, "kernel": kernel, "kernel_ms": ms_kernel, "samples_per_launch": B,
            "algorithmic_flops_per_sample": flops_alg, "executed_flops_per_sample": flops_exe,
            "peak_source": "torch.matmul TF32 8192^3 sustained, measured before warm-up in this run, / 3 (3xTF32 split: "
                           "fp32-equivalent).  `achieved` counts the survey's algorithmic flops (pruned Jacobian form), "
                           "`achieved_executed` the flops the contracted-form GEMMs execute on the zero-padded widths "
                           "(gpflow_b200/roofline.py) -- the tensor pipe does 3x that in TF32 products",
            "slab_samples": int(lib.gpflow_tc_slab_samples(ctypes.byref(desc), B, _lib.OP_TRAIN_STEP)),
        }
        launches = tc_launches + steps        # + gpflow_adam_step
    else:
        peak = peaks["fp32_fma_tflops"]
        peak_hot = ops.fp32_fma_peak_tflops()
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic_bytes_per_launch.json")
        if os.path.exists(tp) and euler is None:
            try:
                traffic = json.load(open(tp))[name]["dram_bytes_per_sample"] * B   # per launch, like `achieved`
            except Exception:
                traffic = None
        achieved_exe, flops_exe = achieved, flops_alg
        if "sf2r" in kernel:   # reverse-mode grad psi: fewer flops than the forward-mode count of the survey
            from gpflow_b200.roofline import executed_flops_per_sample_sf2r
            flops_exe = executed_flops_per_sample_sf2r(wl["hidden"][0], wl["hidden"][1], wl["n_steps"])
            achieved_exe = achieved * flops_exe / flops_alg
        roofline = {
            "bound": "fp32_fma", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
            "frac_nominal": achieved / NOMINAL_FP32_TFLOPS, "peak_nominal": NOMINAL_FP32_TFLOPS,
            "peak_hot_after_run": peak_hot, "peak_sm_mhz": peaks["fp32_fma_sm_mhz"],
            "achieved_executed": achieved_exe, "frac_executed": achieved_exe / peak,
            "executed_flops_per_sample": flops_exe, "traffic": traffic,
            "kernel": kernel, "kernel_ms": ms_kernel, "algorithmic_flops_per_sample": flops_alg, "samples_per_launch": B,
            "peak_source": "FFMA burn micro-kernel (gpflow_fp32_fma_burn) measured ONCE before warm-up in this run, SM clock "
                           "beside it; MEASURED_PEAKS.json records no FP32 CUDA-core peak.  Nominal 148 SM x 128 lanes x "
                           "2 x 1.965 GHz = 74.5 TFLOP/s; `peak_hot_after_run` is the same burn on the hot chip",
            "hbm": {"algorithmic_GBps": 12 * d * B / (ms_kernel * 1e-3) / 1e9,
                    "traffic_GBps": (traffic / (ms_kernel * 1e-3) / 1e9) if traffic else None,
                    "peak_GBps": peaks.get("hbm_gbs_measured"), "peak_kind": "of measured (MEASURED_PEAKS.json copy bandwidth)",
                    "note": "algorithmic bytes = x_data + x_map + y (12 d B/sample)"},
        }
        launches = steps * (ops.launch_count(cfg) + 1)   # + gpflow_adam_step
    cpu_baseline = None
    if with_cpu_baseline and world == 1:
        threads = os.cpu_count() or 1
        rate, sec = cpu_oracle_rate(wl, args.cpu_baseline_sample, 2, 1, threads)
        cpu_baseline = {"value": rate, "unit": "samples/s", "cores": threads, "kind": "port",
                        "sample": "%d samples/step x 2 steps of the same synthetic workload (oracle port, torch CPU "
                                  "autograd, all host threads; calibration against the unmodified reference: BASELINE.md "
                                  "section 2)" % args.cpu_baseline_sample}
    ws_mb = torch.cuda.max_memory_allocated(dev) >> 20
    return {
        "metric": "gp_flow_train_step_samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": world,
        "steps": steps, "warmup": warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak" if weak else "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": name + ": " + wl["desc"], "global_batch": B_global, "per_gpu_batch": B,
                   "d": d, "nb_div_blocks": wl["nb"], "hidden": list(wl["hidden"]), "nb_layers_gp": wl["n_steps"],
                   "euler_coef": euler, "parallelism": "dp%d (batch shards, one all-reduce of %d floats)" % (
                       world, cfg.n_params + 3),
                   "l2": ("inputs larger than L2: rotating through %d distinct resident batches of %.0f MB" % (
                       n_buf, bytes_per_batch / 1e6)) if bytes_per_batch * n_buf > 126e6 else
                         ("the step's working set (slab activations; %d MB allocated) is larger than L2; %d input batches"
                          % (ws_mb, n_buf)),
                   "step": "fused fwd+OT+bwd (%s), finalize, all-reduce, fused Adam (gpflow_adam_step)" % kernel},
        "clocks": clk, "e2e": e2e, "gpu_launches": launches,
        "roofline": roofline, "cpu_baseline": cpu_baseline,
        "loss": float(loss), "reg": float(reg),
        **({"per_rank": per_rank} if per_rank is not None else {}),
    }


if __name__ == "__main__":
    main()
