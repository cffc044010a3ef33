#!/usr/bin/env python
"""bench.py -- the driver's benchmark contract for the examples/mnist DDP training step.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch 64] [--impl reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the hot path over one batch of synthetic 1x28x28 images per rank:
zero_grad, forward, NLL loss, backward, (flat-bucket all_reduce when N>1), fused SGD+momentum
(mnist.py:39-43).  Workload at N=1: BASELINE.json configs[1] (Net on 1xB200, batch 64).
Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for every field.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "MNIST-CNN DDP images/sec"
UNIT = "images/s"
FLOPS_PER_IMAGE = 13_182_000          # 2 x 6,591,000 MAC, dense, fwd+bwd
PARAM_BYTES = 1_724_320
HBM_BYTES_PER_IMAGE = 80_608          # SURVEY.md 8(d) ideal-fusion model
RING_DEVICE = 1024                    # distinct device-resident batches (205 MB > 126 MB L2)
RING_HOST = 256                       # distinct pinned host batches for the e2e leg
SETTLE_S = 0.3                        # untimed load phase in front of each timed region (clock samples under load)

# Algorithmic work of each kernel per image (dense MACs; bytes = tensors it must read+write once,
# weights counted once per launch) -- DESIGN.md "Kernels".
KERNEL_WORK = {
    # name: (MAC per image, bytes per image, bytes per launch)
    "conv1_relu_pool_fwd": (288_000, 3136 + 11520 + 2880, 2080),
    "conv2_relu_pool_fwd": (1_600_000, 11520 + 3200 + 800, 100_200),
    "fc1 forward gemm": (400_000, 3200 + 2000, 1_602_000),
    "fc2/log_softmax/nll head": (10_000, 2000 + 2000 + 80 + 8, 20_040),
    "fc backward (fc1 wgrad + fc1 dgrad + fc2 wgrad)": (805_000, 2000 * 2 + 3200 + 3200 + 2000 + 40, 3_220_000),
    "conv2 backward (wgrad + dgrad)": (3_200_000, 3200 + 3200 + 800 + 11520 + 11520, 200_200),
    "conv1 wgrad": (288_000, 3136 + 11520 + 11520 + 2880, 2080),
    # large batches (>= 1024): conv2 on tcgen05, forward / data gradient / weight gradient as separate launches
    "conv2_relu_pool_fwd (tcgen05 3xTF32 shifted-window GEMM)": (1_600_000, 11520 + 3200 + 800, 100_200 + 256_000),
    "conv2 dgrad (tcgen05 3xTF32 GEMM + col2im)": (1_600_000, 3200 + 3200 + 800 + 11520, 229_376),
    "conv2 wgrad (tcgen05 3xTF32, all four hi/lo products per MMA)": (1_600_000, 3200 + 3200 + 800 + 11520,
                                                                     29 * 25088 * 4),
    "conv2 wgrad": (1_600_000, 3200 + 3200 + 800 + 11520, 32 * 25088 * 4),
    "conv2 dgrad": (1_600_000, 3200 + 3200 + 800 + 11520, 100_000),
    "conv2 weight split (hi/lo)": (0, 0, 100_000 + 256_000),
    "conv2 dgrad weight split (hi/lo)": (0, 0, 100_000 + 229_376),
    "conv2 weight split (hi/lo; forward + dgrad operands)": (0, 0, 100_000 + 256_000 + 229_376),
    "fc1 forward gemm (tcgen05 3xTF32)": (400_000, 3200 + 2000, 1_602_000),
    "fc1 dgrad gemm (tcgen05 3xTF32)": (400_000, 2000 + 3200, 1_600_000),
    "fc1 wgrad gemm (tcgen05 3xTF32)": (400_000, 2000 + 3200, 1_602_000),
    "fc1 wgrad + dgrad gemms (tcgen05 3xTF32, one launch)": (800_000, 2000 + 3200 + 3200, 3_202_000),
    "fc1 operand split (p2 hi/lo)": (0, 3200 + 2 * 3328, 0),
    "fc1 operand split (W1 hi/lo)": (0, 0, 3 * 1_600_000),
    "fc1 operand split (dz1 hi/lo)": (0, 3 * 2000, 0),
    "sgd momentum": (0, 0, 5 * PARAM_BYTES),
    "grad finalize": (0, 0, 2 * 31_080 * 4),
}


def load_traffic(batch):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of each kernel, from the committed ncu --set full
    capture of this workload (profiles/r01_traffic.json, written by tools/ncu_traffic.py); None if absent."""
    try:
        with open(os.path.join(ROOT, "profiles", "r01_traffic.json")) as f:
            return json.load(f).get(f"batch{batch}", {})
    except Exception:
        return {}


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            p = json.load(f)
        return {"hbm_gbs": float(p["hbm_gbs"]), "bf16_tflops": float(p["bf16_tflops"]),
                "bf16_tflops_sustained": float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), "source": "measured"}
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons for one GPU while a timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def window(self, t0, t1):
        return [r for (t, r) in self.rows if t0 <= t <= t1 + 0.06]

    def under_load(self, t0, t1):
        """Clock summary of the timed region [t0, t1]; a region shorter than two sampling periods is judged by the
        samples of the untimed load phase that runs directly before it (settle()), which says so in `note`."""
        rows, note = self.window(t0, t1), None
        if len(rows) < 2:
            rows = self.window(t0 - SETTLE_S, t1)
            note = (f"timed region of {1e3 * (t1 - t0):.1f} ms is shorter than the 50 ms sampling period: samples of the "
                    f"{SETTLE_S} s untimed load phase directly before it are included")
        c = ClockSampler.summarise(rows)
        if c is not None and note:
            c["note"] = note
        return c

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    @staticmethod
    def summarise(lines):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in lines:
            parts = [p.strip() for p in ln.split(",")]
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except Exception:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return None
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference_run(batch, steps, warmup, budget_s, threads=None):
    """The reference's train() body on the host cores (oracle/mnist_oracle.TorchCpuTrainer: the same
    torch CPU ops, autograd and optim.SGD the reference payload uses).  Returns a dict."""
    import torch
    from oracle import mnist_oracle as mo
    if threads:
        torch.set_num_threads(threads)
    cores = torch.get_num_threads()
    tr = mo.TorchCpuTrainer(seed=1)
    ring = [mo.synthetic_batch(batch, 1234 + i) for i in range(16)]
    for i in range(warmup):
        tr.step(*ring[i % 16])
    t0 = time.perf_counter()
    done = 0
    while done < steps:
        tr.step(*ring[done % 16])
        done += 1
        if time.perf_counter() - t0 > budget_s and done >= min(steps, REFERENCE_MIN_STEPS):
            break
    dt = time.perf_counter() - t0
    return {"value": batch * done / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{done} training steps of batch {batch} (reference train() body on torch {torch.__version__} "
                      f"CPU ops, {cores} threads, {os.cpu_count()} host cpus), {dt:.1f} s",
            "ms_per_step": 1e3 * dt / done, "steps": done}


def cpu_reference_run_ddp(batch, steps, warmup, budget_s, world, hard_timeout_s=150.0):
    """The reference's CPU/gloo configuration (BASELINE configs[0] generalised to `world` ranks): `world` fresh
    local processes, DistributedDataParallel on gloo, host cores split evenly between the ranks.  The processes get
    a clean environment and their own tcp:// store (oracle.mnist_oracle.run_ddp_cpu), and are killed after
    `hard_timeout_s`, so this arm also works -- and stays bounded -- when bench.py itself runs under torchrun."""
    import torch
    from oracle import mnist_oracle as mo
    cores = os.cpu_count() or 1
    threads = max(1, cores // world)
    r = mo.run_ddp_cpu(world, batch, steps, warmup, budget_s, threads, hard_timeout_s)
    return {"value": r["value"], "unit": UNIT, "cores": threads * world, "kind": "port",
            "sample": f"{r['steps']} DDP training steps, {world} gloo ranks x batch {batch} x {threads} threads "
                      f"(reference train() body + DistributedDataParallel on torch {torch.__version__} CPU ops, "
                      f"{cores} host cpus), {r['seconds']:.1f} s",
            "ms_per_step": r["ms_per_step"], "steps": r["steps"]}


REFERENCE_MIN_STEPS = 200   # a 20-step CPU sample swings by +-15 %: the arm runs for about --ref-budget seconds instead
REFERENCE_STEPS = 1_000_000  # never reached: the time budget ends the run


def run_reference_arm(args, rank):
    """bench.py --impl reference: the reference's own CPU implementation of the step (torch CPU ops, autograd,
    optim.SGD, gloo DistributedDataParallel at N>1) on the box's host cores.  Under torchrun only rank 0 works; the
    other ranks return at once (exit 0)."""
    if rank != 0:
        return
    steps = REFERENCE_STEPS
    warmup = max(args.warmup, 3)
    note = ""
    r = None
    if args.gpus > 1:
        try:
            r = cpu_reference_run_ddp(args.batch, steps, warmup, budget_s=args.ref_budget, world=args.gpus)
            how = f"reference CPU path: {args.gpus} gloo ranks with DistributedDataParallel, per-rank batch {args.batch}"
        except Exception as e:   # bounded by the watchdog; fall back to the single-process arm rather than no number
            note = f"; gloo DDP arm failed ({e!r}), single-process CPU run reported instead"
    if r is None:
        import torch
        torch.set_num_threads(os.cpu_count() or 1)   # torchrun exports OMP_NUM_THREADS=1 to its workers
        r = cpu_reference_run(args.batch, steps, warmup, budget_s=args.ref_budget)
        how = "reference CPU path, single process"
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": r["steps"], "warmup": warmup, "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"examples/mnist Net DDP train step (BASELINE configs[{1 if args.gpus == 1 else 2}]), "
                                   f"per-rank batch {args.batch}, synthetic 1x28x28", "global_batch": args.gpus * args.batch,
                       "how": how + f"; about {args.ref_budget:g} s of timed steps (never fewer than {REFERENCE_MIN_STEPS}), whatever --steps says" + note},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def load_compute_peaks():
    """fp32 FMA and TF32 tensor peaks of this GPU: profiles/r02_peaks_fp32_tf32.json, measured on the B200 box by
    tools/measure_peaks.py the way MEASURED_PEAKS.json was made; datasheet figures if that file is missing."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_peaks_fp32_tf32.json")) as f:
            p = json.load(f)
        return {"fp32_tflops": float(p["fp32_ffma_tflops"]), "tf32_tflops": float(p["tf32_tflops"]),
                "source": "measured (profiles/r02_peaks_fp32_tf32.json, tools/measure_peaks.py)"}
    except Exception:
        return {"fp32_tflops": 80.0, "tf32_tflops": 1100.0, "source": "fallback (B200 datasheet: 80 fp32, 1100 dense TF32)"}


def kernel_roofline(name, us, B, peaks, cpeaks, traffic):
    """Roofline of one kernel launch.  The compute ceiling is the one the kernel's arithmetic can reach: the measured
    fp32 FMA peak for the SIMT (FFMA2) kernels, the measured TF32 tensor peak / 3 for the tcgen05 3xTF32 kernels
    (three MMAs per algorithmic product; the conv2 weight gradient issues four), none for pure data movement.  The
    bound is whichever of compute time and HBM time is longer for the ALGORITHMIC work."""
    mac, bpi, bpl = KERNEL_WORK.get(name, (0, 0, 0))
    flops, nbytes, t = 2.0 * mac * B, float(bpi * B + bpl), us * 1e-6
    tc = "tcgen05" in name
    cpeak = (cpeaks["tf32_tflops"] / 3.0) if tc else cpeaks["fp32_tflops"]
    t_c, t_m = flops / (cpeak * 1e12), nbytes / (peaks["hbm_gbs"] * 1e9)
    if t_c >= t_m and flops > 0:
        roof = {"bound": "tensor" if tc else "fp32", "achieved": flops / t / 1e12, "peak": cpeak, "unit": "TFLOP/s",
                "peak_source": cpeaks["source"] + ("; TF32 dense peak / 3 (3xTF32 split)" if tc else "; packed fp32 FMA loop")}
    else:
        roof = {"bound": "hbm", "achieved": nbytes / t / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "peak_source": peaks["source"] + " (MEASURED_PEAKS.json copy bandwidth)"}
    roof["frac"] = roof["achieved"] / roof["peak"]
    roof.update({"traffic": traffic.get(name), "kernel": name, "kernel_us": us, "algorithmic_flops_per_launch": flops,
                 "algorithmic_bytes_per_launch": nbytes, "hbm_frac": nbytes / t / 1e9 / peaks["hbm_gbs"],
                 "compute_frac": (flops / t / 1e12 / cpeak) if flops else 0.0})
    return roof


class Bench:
    """One model / optimizer / process group; run_config() measures one batch size."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist

        import pytorch_operator_b200 as pkg
        self.torch, self.dist, self.pkg, self.args = torch, dist, pkg, args
        self.rank = int(os.environ.get("RANK", 0))
        self.world = int(os.environ.get("WORLD_SIZE", 1))
        local_rank = int(os.environ.get("LOCAL_RANK", self.rank))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a GPU: the CUDA path has no CPU fallback")
        if self.world != args.gpus:
            raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={self.world}: launch with torch.distributed.run")
        ndev = torch.cuda.device_count()
        self.dev = torch.device("cuda", local_rank % ndev)
        torch.cuda.set_device(self.dev)
        if self.world > 1:
            if os.environ.get("NCCL_DEBUG"):  # keep stdout clean for the single JSON line
                os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
            dist.init_process_group("nccl", device_id=self.dev)
        self.lib = pkg._cabi.lib()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        self.sampler = ClockSampler(self.dev.index if not vis else int(vis.split(",")[self.dev.index]))
        self.peaks, self.cpeaks = load_peaks(), load_compute_peaks()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, ms):
        if self.world == 1:
            return ms
        t = self.torch.tensor([ms], device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t)

    # ---- parity prologue (N > 1): put the multi-GPU checks on the record before anything is timed ----
    def parity_prologue(self):
        """tests/ddp_worker.py's checks, in this process group: the reference's ws=2 gloo fixture (N=2), fused sum ==
        rank-ordered sum bit for bit, replicas bit-identical after every step (eager and CUDA-graph steps), smoke-dist
        send/recv + all_reduce bit-exact.  Returns the parity block; any failure on any rank ends the run non-zero."""
        torch, dist = self.torch, self.dist
        from tests import ddp_worker as dw
        probe = self.pkg.DistributedDataParallel(self.pkg.Net().to(self.dev))  # which exchange path does this box give?
        fused = bool(probe.fused)
        why = probe.fused_error
        probe.close()
        out = {"exchange": "fused peer-window kernels" if fused else f"nccl all_reduce ({why})"}

        def run(key, fn):
            try:
                fn()
                ok = 1
            except Exception as e:  # AssertionError from a check, or a CUDA/NCCL error
                ok = 0
                print(f"rank {self.rank}: parity check {key} failed: {e!r}", file=sys.stderr, flush=True)
            flag = torch.tensor([ok], device=self.dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            out[key] = "pass" if int(flag) == 1 else "FAIL"

        if self.world == 2:
            run("ws2_fixture", lambda: dw.fixture_run(self.dev, fused=fused, graph=True))
        else:
            out["ws2_fixture"] = "n/a (the reference fixture is a world_size-2 run)"
        run("rank_ordered_sum", lambda: dw.property_run(self.dev, fused=fused, graph=True, steps=5))
        out["replicas_identical"] = out["rank_ordered_sum"]   # one run checks both after every step
        run("smoke_dist", lambda: dw.smoke_dist_run(sizes=(4, 431296), iters=3))
        out["ok"] = all(out[k] == "pass" or out[k].startswith("n/a") for k in ("ws2_fixture", "rank_ordered_sum", "smoke_dist"))
        return out

    def build_model(self):
        torch, pkg = self.torch, self.pkg
        torch.manual_seed(1)
        self.net = pkg.Net().to(self.dev)
        self.model = pkg.DistributedDataParallel(self.net) if self.world > 1 else self.net
        self.opt = pkg.SGD(self.model.parameters(), lr=0.01, momentum=0.5)

    # ---- one workload ----
    def settle(self, step, sync=None):
        """Untimed load phase in front of a timed region: `step(i)` over and over for SETTLE_S seconds (every rank the
        same number of times: the step may contain a cross-rank exchange)."""
        torch = self.torch
        sync = sync or torch.cuda.synchronize
        if SETTLE_S <= 0.0:
            return 0
        n, t0 = 0, time.time()
        while True:
            for _ in range(64):
                step(n)
                n += 1
            sync()
            go = torch.tensor([1.0 if time.time() - t0 < SETTLE_S else 0.0], device=self.dev)
            if self.world > 1:
                self.dist.all_reduce(go, op=self.dist.ReduceOp.MIN)
            if float(go) == 0.0:
                return n

    def run_config(self, B, K, W, profile_iters):
        torch, pkg, lib, dev, world, rank = self.torch, self.pkg, self.lib, self.dev, self.world, self.rank
        model, net, opt = self.model, self.net, self.opt
        # synthetic data: packed [x fp32 | target int64] records, one record per batch
        xb, tb = B * 784 * 4, B * 8
        rec = xb + tb
        nring = max(8, min(RING_DEVICE, (256 << 20) // rec))
        g = torch.Generator(device=dev).manual_seed(1234 + rank)
        ring = torch.empty(nring, rec, dtype=torch.uint8, device=dev)
        ring[:, :xb].view(torch.float32).normal_(generator=g)
        ring[:, xb:].view(torch.int64).copy_(torch.randint(0, 10, (nring, B), generator=g, device=dev))
        nhost = max(4, min(RING_HOST, (64 << 20) // rec))
        host_ring = torch.empty(nhost, rec, dtype=torch.uint8).pin_memory()
        host_ring.copy_(ring[:nhost])
        static = torch.zeros(rec, dtype=torch.uint8, device=dev)
        sx = static[:xb].view(torch.float32).view(B, 1, 28, 28)
        st = static[xb:].view(torch.int64)

        # first step eager (initialises momentum), count kernels per step, capture the graph
        self.barrier()
        static.copy_(ring[0])
        model.train_step(sx, st, opt)
        torch.cuda.synchronize()
        c0 = lib.b200mnist_launch_count()
        model.train_step(sx, st, opt)
        torch.cuda.synchronize()
        kernels_per_step = int(lib.b200mnist_launch_count() - c0)
        # leg 1 (value): the batches are resident in HBM (the device ring) and every graph replay reads the next slot in
        # place through the ring's device-side cursor (InputRing) -- no per-step copy, no host-side pointer.  On the
        # NCCL DDP path, which has no ring form of the step, each batch is copied into the static buffer first.
        in_place = world == 1 or bool(getattr(model, "fused", False))
        in_ring = pkg.InputRing(B, records=ring) if in_place else None
        # with the ring a graph may hold several consecutive steps (a graph boundary costs ~1.5 us of GPU time at
        # B=64): the K timed steps are K / unroll replays
        unroll = (4 if K % 4 == 0 else 2 if K % 2 == 0 else 1) if in_place else 1
        graph = pkg.StepGraph(model, opt, B, x=sx, target=st, ring=in_ring, unroll=unroll)
        def one_replay(i):
            if not in_place:
                static.copy_(ring[i % nring], non_blocking=True)
            graph.replay()

        self.settle(one_replay)   # untimed: the same load for SETTLE_S, so that clock samples under load exist
        for i in range((W + unroll - 1) // unroll):
            one_replay(i)
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t_start = time.time()
        e0.record()
        for i in range(K // unroll):
            if not in_place:
                static.copy_(ring[i % nring], non_blocking=True)
            graph.replay()
        e1.record()
        self.barrier()
        t_end = time.time()
        ms_total = self.max_over_ranks(e0.elapsed_time(e1))
        clocks = self.sampler.under_load(t_start, t_end)
        loss_final = float(graph.loss)

        # leg 2: end to end from pinned host memory, loss read back every step.  Public API:
        # StepPipeline.submit(pinned record) = H2D on a copy stream into a slot of a small device ring + the slot's sequence
        # number behind it, graph replay (the step waits for its slot on the device), {step, loss} stored into pinned host
        # memory by the step's last kernel -- all inside the timed region, every step.
        pipe = pkg.StepPipeline(model, opt, B, unroll=unroll)
        self.settle(lambda i: pipe.submit(host_ring[i % nhost]), sync=pipe.drain)
        for i in range((W + unroll - 1) // unroll * unroll):
            pipe.submit(host_ring[i % nhost])
        pipe.drain()
        self.barrier()
        t2_start = time.time()
        e0.record()
        for i in range(K):
            pipe.submit(host_ring[i % nhost])
        pipe.flush()       # steps still waiting for their group of `unroll` go out now: all K steps precede e1
        e1.record()
        pipe.drain()
        if pipe.steps_done() != pipe.steps:
            raise SystemExit("the e2e leg did not run every submitted step")
        self.barrier()
        t2_end = time.time()
        ms_e2e = self.max_over_ranks(e0.elapsed_time(e1))
        clocks_e2e = self.sampler.under_load(t2_start, t2_end)
        if not bool(torch.isfinite(pipe.loss_host).all()):
            raise SystemExit("non-finite loss in the timed region")

        # fused exchange: where the exchange sat inside the LAST timed step (GPU globaltimer stamps its kernels leave
        # in the flag page; microseconds after the exchange CTAs started), max over ranks
        timeline = None
        if world > 1 and getattr(model, "fused", False):
            tel = model.exchange_telemetry()
            timeline = {k: round(self.max_over_ranks(v), 2) for k, v in tel.items()}

        net.release_graphs()
        del graph, pipe
        out = {"B": B, "K": K, "W": W, "rec": rec, "nring": nring, "ms_total": ms_total, "ms_e2e": ms_e2e,
               "kernels_per_step": kernels_per_step, "clocks": clocks, "clocks_e2e": clocks_e2e,
               "loss_final": loss_final, "timeline": timeline, "unroll": unroll}
        if rank != 0:
            return out

        # per-kernel durations (rank 0): CUDA events on the launching stream after every kernel of this rank's step
        # (the collective is timed separately).  Event-to-event times carry the cost of the event itself and lose
        # the programmatic-dependent-launch overlap, so they are scaled to add up to the same eager step timed with
        # one event pair around it (both runs queue behind a delay kernel: no host gaps).
        import ctypes
        names_buf = (ctypes.c_char_p * 64)()
        ms_buf = (ctypes.c_float * 64)()
        stream = torch.cuda.current_stream().cuda_stream
        sums, counts = {}, {}
        saved = [t.clone() for t in (net.flat_params, net.flat_grads, net.flat_momentum)]
        ddp_saved = net._ddp
        object.__setattr__(net, "_ddp", None)
        plain_us = 0.0
        for it in range(profile_iters + 3):
            static.copy_(ring[it % nring], non_blocking=True)
            pkg._cabi.check(lib.b200mnist_profile_begin(stream, 400), "profile_begin")
            net.train_step(sx, st, opt)
            n = lib.b200mnist_profile_end(ms_buf, names_buf, 64, stream)
            if n < 0:
                raise SystemExit("profile_end failed: " + pkg._cabi.last_error())
            lib.b200mnist_delay(stream, 400)
            e0.record()
            net.train_step(sx, st, opt)
            e1.record()
            torch.cuda.synchronize()
            if it < 3:
                continue
            plain_us += e0.elapsed_time(e1) * 1e3 / profile_iters
            for i in range(n):
                nm = names_buf[i].decode()
                sums[nm] = sums.get(nm, 0.0) + ms_buf[i]
                counts[nm] = counts.get(nm, 0) + 1
        object.__setattr__(net, "_ddp", ddp_saved)
        for t, sv in zip((net.flat_params, net.flat_grads, net.flat_momentum), saved):
            t.copy_(sv)
        torch.cuda.synchronize()
        raw_us = {k: 1e3 * sums[k] / counts[k] for k in sums}
        raw_sum = sum(raw_us.values())
        scale = min(1.0, plain_us / raw_sum) if raw_sum > 0 else 1.0
        out.update({"kern_us": {k: v * scale for k, v in raw_us.items()}, "kern_us_raw_sum": raw_sum,
                    "eager_step_us": plain_us, "event_scale": scale})
        return out

    def report(self, r):
        """JSON fields of one measured config."""
        B, K, world = r["B"], r["K"], self.world
        step_us = 1e3 * r["ms_total"] / K
        kern_us = r["kern_us"]
        top = max(kern_us, key=kern_us.get)
        traffic = load_traffic(B)
        roof = kernel_roofline(top, kern_us[top], B, self.peaks, self.cpeaks, traffic)
        roof["note"] = ("dominant kernel of the step by duration; kernel_us = CUDA-event time of the kernel on its launching "
                        "stream, scaled so that the kernels add up to the eager step timed with one event pair")
        step_flops, step_bytes = FLOPS_PER_IMAGE * B, HBM_BYTES_PER_IMAGE * B + 8 * PARAM_BYTES
        t = step_us * 1e-6
        simt_share = sum(v for k, v in kern_us.items() if "tcgen05" not in k) / max(sum(kern_us.values()), 1e-9)
        return {
            "ms_per_step": r["ms_total"] / K, "img_s": world * B * K / (r["ms_total"] * 1e-3),
            "e2e": {"value": world * B * K / (r["ms_e2e"] * 1e-3), "unit": UNIT, "h2d_bytes_per_step": r["rec"],
                    "d2h_bytes_per_step": 4, "ms_per_step": r["ms_e2e"] / K},
            "kernels_per_step": r["kernels_per_step"], "roofline": roof,
            "step_roofline": {"flops_per_step": step_flops, "hbm_bytes_per_step": step_bytes,
                              "tflops": step_flops / t / 1e12, "hbm_gbs": step_bytes / t / 1e9,
                              "frac_of_hbm_peak": step_bytes / t / 1e9 / self.peaks["hbm_gbs"],
                              "frac_of_fp32_peak": step_flops / t / 1e12 / self.cpeaks["fp32_tflops"],
                              "frac_of_tf32_peak_over_3": step_flops / t / 1e12 / (self.cpeaks["tf32_tflops"] / 3.0),
                              "simt_share_of_kernel_time": simt_share},
            "kernel_us": {k: round(v, 3) for k, v in sorted(kern_us.items(), key=lambda kv: -kv[1])},
            "kernel_us_sum": round(sum(kern_us.values()), 3), "kernel_us_raw_event_sum": round(r["kern_us_raw_sum"], 3),
            "eager_step_us": round(r["eager_step_us"], 3), "clocks": r["clocks"], "clocks_e2e": r["clocks_e2e"],
            "final_loss": r["loss_final"], "exchange_timeline_us": r["timeline"],
        }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20000)
    ap.add_argument("--warmup", type=int, default=200)
    ap.add_argument("--batch", type=int, default=64, help="per-rank batch (BASELINE configs[1]: 64; config 5: 4096)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ref-budget", type=float, default=10.0, help="--impl reference: seconds of timed CPU work")
    ap.add_argument("--profile-iters", type=int, default=30)
    ap.add_argument("--no-config5", action="store_true", help="skip the batch-4096 leg (BASELINE configs[4])")
    ap.add_argument("--no-parity", action="store_true", help="skip the multi-GPU parity prologue (N > 1)")
    ap.add_argument("--settle", type=float, default=0.3,
                    help="seconds of untimed load in front of each timed region (0 under ncu, where every launch is replayed)")
    args = ap.parse_args()
    globals()["SETTLE_S"] = max(0.0, args.settle)
    rank = int(os.environ.get("RANK", 0))

    if args.impl == "reference":
        run_reference_arm(args, rank)
        return

    b = Bench(args)
    torch, dist, world = b.torch, b.dist, b.world
    W, K, B = max(args.warmup, 3), args.steps, args.batch

    parity = None
    if world > 1 and not args.no_parity:
        parity = b.parity_prologue()
        if not parity["ok"]:
            if rank == 0:
                print(json.dumps({"metric": METRIC, "value": None, "n_gpus": world, "parity": parity,
                                  "error": "multi-GPU parity check failed; nothing was timed"}), flush=True)
            dist.barrier()
            dist.destroy_process_group()
            sys.exit(1)

    b.build_model()
    main_r = b.run_config(B, K, W, args.profile_iters)
    c5 = None
    if not args.no_config5 and B != 4096:
        # BASELINE configs[4]: 4096 images per rank, the tcgen05 kernels (>= 10 warm-up + >= 100 steps)
        c5 = b.run_config(4096, max(100, min(K, 300)), max(10, min(W, 30)), min(args.profile_iters, 10))

    # gradient all_reduce bus bandwidth (the second half of BASELINE.json's metric), all ranks
    busbw = None
    if world > 1:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        buf = torch.zeros(b.net.flat_grads.numel(), dtype=torch.float32, device=b.dev)
        for _ in range(5):
            dist.all_reduce(buf)
        b.barrier()
        e0.record()
        for _ in range(50):
            dist.all_reduce(buf)
        e1.record()
        torch.cuda.synchronize()
        ar_ms = b.max_over_ranks(e0.elapsed_time(e1) / 50)
        nbytes_ar = buf.numel() * 4
        busbw = {"bytes": nbytes_ar, "nccl_time_us": ar_ms * 1e3,
                 "nccl_busbw_GBps": nbytes_ar / (ar_ms * 1e-3) * 2 * (world - 1) / world / 1e9}
    b.sampler.stop()
    fused = bool(getattr(b.model, "fused", False))

    def teardown():
        if world > 1:
            b.net.release_graphs()  # a graph holding NCCL kernels would make destroy_process_group() block
            b.model.close()
            dist.barrier()
            dist.destroy_process_group()

    if rank != 0:
        teardown()
        return

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cpu = cpu_reference_run(B, 2000, 20, budget_s=12.0)
        cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}

    m = b.report(main_r)
    line = {
        "metric": METRIC, "value": m["img_s"], "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": m["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"examples/mnist Net DDP train step (BASELINE configs[{1 if world == 1 else 2}]), "
                               f"per-rank batch {B}, synthetic 1x28x28", "global_batch": world * B,
                   "parallelism": f"dp{world}", "l2": f"inputs cycle through {main_r['nring']} distinct device-resident "
                                                      f"batches ({main_r['nring'] * main_r['rec'] / 1e6:.0f} MB > 126 MB L2), "
                                                      "read in place through a device-side cursor",
                   "cuda_graph": True, "steps_per_graph": main_r["unroll"], "lr": 0.01, "momentum": 0.5,
                   "untimed_load_before_each_timed_region_s": SETTLE_S},
        "e2e": dict(m["e2e"], how="StepPipeline.submit(pinned record): H2D on a copy stream into a slot of a small device "
                                  "ring with the slot's sequence number written behind it, graph replay (the step waits for "
                                  "its slot on the device), {step, loss} stored by the last kernel straight into pinned "
                                  "host memory (zero-copy D2H) -- every step."),
        "gpu_launches": m["kernels_per_step"] * K, "kernels_per_step": m["kernels_per_step"],
        "clocks": m["clocks"], "clocks_e2e": m["clocks_e2e"], "roofline": m["roofline"],
        "step_roofline": m["step_roofline"], "kernel_us": m["kernel_us"], "kernel_us_sum": m["kernel_us_sum"],
        "kernel_us_raw_event_sum": m["kernel_us_raw_event_sum"], "eager_step_us": m["eager_step_us"],
        "ddp_exchange": (None if world == 1 else ("fused into the step's kernels over NVLink peer windows (CUDA IPC): "
                                                  "two-shot fc1 slice reduce+SGD+push by extra CTAs of the conv1 "
                                                  "weight-gradient launch, remainder pushed + summed by the last kernel"
                                                  if fused else "NCCL all_reduce + SGD kernel")),
        "exchange_timeline_us": m["exchange_timeline_us"], "grad_allreduce": busbw, "parity": parity,
        "cpu_baseline": cpu, "final_loss": m["final_loss"],
    }
    if c5 is not None:
        c = b.report(c5)
        line["config5"] = {"workload": "examples/mnist Net DDP train step (BASELINE configs[4]), per-rank batch 4096, "
                                       "tcgen05 conv2 / fc1 kernels", "global_batch": world * 4096, "steps": c5["K"],
                           "warmup": c5["W"], **{k: c[k] for k in ("ms_per_step", "img_s", "e2e", "kernels_per_step",
                                                                    "roofline", "step_roofline", "kernel_us",
                                                                    "kernel_us_sum", "eager_step_us", "clocks",
                                                                    "final_loss", "exchange_timeline_us")}}
    print(json.dumps(line), flush=True)
    teardown()


if __name__ == "__main__":
    main()
