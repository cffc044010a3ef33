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
        if time.perf_counter() - t0 > budget_s:
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


REFERENCE_MIN_STEPS = 200   # a 20-step CPU sample swings by +-15 %; the arm is time-budgeted instead


def run_reference_arm(args, rank):
    """bench.py --impl reference: the reference's own CPU implementation of the step (torch CPU ops, autograd,
    optim.SGD, gloo DistributedDataParallel at N>1) on the box's host cores.  Under torchrun only rank 0 works; the
    other ranks return at once (exit 0)."""
    if rank != 0:
        return
    steps = max(args.steps, REFERENCE_MIN_STEPS)
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
                       "how": how + f"; at least {REFERENCE_MIN_STEPS} timed steps, capped by a {args.ref_budget:g} s budget" + note},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20000)
    ap.add_argument("--warmup", type=int, default=200)
    ap.add_argument("--batch", type=int, default=64, help="per-rank batch (BASELINE configs[1]: 64; config 5: 4096)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ref-budget", type=float, default=20.0, help="--impl reference: seconds of timed CPU work")
    ap.add_argument("--profile-iters", type=int, default=30)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", rank))

    if args.impl == "reference":
        run_reference_arm(args, rank)
        return

    import torch
    import torch.distributed as dist

    import pytorch_operator_b200 as pkg

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU: the CUDA path has no CPU fallback")
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    W = max(args.warmup, 3)
    K = args.steps
    B = args.batch
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        if os.environ.get("NCCL_DEBUG"):  # keep stdout clean for the single JSON line
            os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    lib = pkg._cabi.lib()

    torch.manual_seed(1)
    model = pkg.Net().to(dev)
    net = model
    if world > 1:
        model = pkg.DistributedDataParallel(model)
    opt = pkg.SGD(model.parameters(), lr=0.01, momentum=0.5)

    # ---- synthetic data: packed [x fp32 | target int64] records, one record per batch ----
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

    # ---- first step eager (initialises momentum), count kernels per step, capture the graph ----
    static.copy_(ring[0])
    model.train_step(sx, st, opt)
    torch.cuda.synchronize()
    c0 = lib.b200mnist_launch_count()
    model.train_step(sx, st, opt)
    torch.cuda.synchronize()
    kernels_per_step = int(lib.b200mnist_launch_count() - c0)
    graph = pkg.StepGraph(model, opt, B, x=sx, target=st)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    sampler = ClockSampler(torch.cuda.current_device() if os.environ.get("CUDA_VISIBLE_DEVICES") is None
                           else int(os.environ["CUDA_VISIBLE_DEVICES"].split(",")[local_rank]))

    # ---- leg 1: device-resident inputs (value) ----
    for i in range(W):
        static.copy_(ring[i % nring], non_blocking=True)
        graph.replay()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_start = time.time()
    e0.record()
    for i in range(K):
        static.copy_(ring[i % nring], non_blocking=True)
        graph.replay()
    e1.record()
    barrier()
    t_end = time.time()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    clocks = ClockSampler.summarise(sampler.window(t_start, t_end))
    loss_final = float(graph.loss)

    # ---- leg 2: end to end from pinned host memory, loss read back every step ----
    # public API: StepPipeline.submit(pinned record) = H2D on a copy stream (double-buffered), graph replay,
    # D2H copy node of the loss into pinned host memory -- all inside the timed region, every step.
    pipe = pkg.StepPipeline(model, opt, B)
    for i in range(W):
        pipe.submit(host_ring[i % nhost])
    pipe.drain()
    barrier()
    t2_start = time.time()
    e0.record()
    for i in range(K):
        pipe.submit(host_ring[i % nhost])
    e1.record()
    pipe.drain()
    barrier()
    t2_end = time.time()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    clocks_e2e = ClockSampler.summarise(sampler.window(t2_start, t2_end))
    sampler.stop()
    if not bool(torch.isfinite(pipe.loss_host).all()):
        raise SystemExit("non-finite loss in the timed region")

    # ---- gradient all_reduce bus bandwidth (the second half of BASELINE.json's metric), all ranks ----
    busbw = None
    if world > 1:
        buf = torch.zeros(net.flat_grads.numel(), dtype=torch.float32, device=dev)
        for _ in range(5):
            dist.all_reduce(buf)
        barrier()
        e0.record()
        for _ in range(50):
            dist.all_reduce(buf)
        e1.record()
        torch.cuda.synchronize()
        ar_ms = max_over_ranks(e0.elapsed_time(e1) / 50)
        nbytes_ar = buf.numel() * 4
        busbw = {"bytes": nbytes_ar, "nccl_time_us": ar_ms * 1e3,
                 "nccl_busbw_GBps": nbytes_ar / (ar_ms * 1e-3) * 2 * (world - 1) / world / 1e9}

    if rank != 0:
        if world > 1:
            net.release_graphs()  # a graph holding NCCL kernels would make destroy_process_group() block
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- per-kernel durations, CUDA events on the launching stream (rank 0) ----
    import ctypes
    names_buf = (ctypes.c_char_p * 64)()
    ms_buf = (ctypes.c_float * 64)()
    stream = torch.cuda.current_stream().cuda_stream
    sums, counts = {}, {}
    pnet, popt = net, opt
    saved = [t.clone() for t in (pnet.flat_params, pnet.flat_grads, pnet.flat_momentum)]
    ddp_saved = pnet._ddp
    object.__setattr__(pnet, "_ddp", None)  # time this rank's kernels; the collective is timed separately
    for it in range(args.profile_iters + 3):
        static.copy_(ring[it % nring], non_blocking=True)
        pkg._cabi.check(lib.b200mnist_profile_begin(stream, 400), "profile_begin")
        pnet.train_step(sx, st, popt)
        n = lib.b200mnist_profile_end(ms_buf, names_buf, 64, stream)
        if n < 0:
            raise SystemExit("profile_end failed: " + pkg._cabi.last_error())
        if it < 3:
            continue
        for i in range(n):
            nm = names_buf[i].decode()
            sums[nm] = sums.get(nm, 0.0) + ms_buf[i]
            counts[nm] = counts.get(nm, 0) + 1
    object.__setattr__(pnet, "_ddp", ddp_saved)
    for t, sv in zip((pnet.flat_params, pnet.flat_grads, pnet.flat_momentum), saved):
        t.copy_(sv)
    kern_us = {k: 1e3 * sums[k] / counts[k] for k in sums}
    peaks = load_peaks()
    top = max(kern_us, key=kern_us.get)
    mac, bpi, bpl = KERNEL_WORK.get(top, (0, 0, 0))
    flops = 2.0 * mac * B
    nbytes = float(bpi * B + bpl)
    t_top = kern_us[top] * 1e-6
    t_tensor, t_hbm = flops / (peaks["bf16_tflops"] * 1e12), nbytes / (peaks["hbm_gbs"] * 1e9)
    if t_tensor >= t_hbm:
        roof = {"bound": "tensor", "achieved": flops / t_top / 1e12, "peak": peaks["bf16_tflops"], "unit": "TFLOP/s"}
    else:
        roof = {"bound": "hbm", "achieved": nbytes / t_top / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s"}
    roof["frac"] = roof["achieved"] / roof["peak"]
    roof.update({"traffic": load_traffic(B).get(top), "kernel": top, "kernel_us": kern_us[top], "peak_source": peaks["source"] +
                 " (MEASURED_PEAKS.json burst bf16 / copy bandwidth)", "algorithmic_flops_per_launch": flops,
                 "algorithmic_bytes_per_launch": nbytes,
                 "note": ("tcgen05 3xTF32 kernel: algorithmic (fp32-equivalent) FLOPs of the dense product against the "
                          "bf16 tensor peak; the MMAs issued are 2-4x that (hi/lo products) at TF32 rate (half of bf16)"
                          if "tcgen05" in top else
                          "fp32 SIMT (FFMA2) kernel -- 1e-5 parity rules out single-pass TF32/bf16 MMA; at this batch the "
                          "kernel is latency/occupancy bound (see profiles/), so both roofline fractions are low by construction")})
    step_us = 1e3 * ms_total / K
    step_flops, step_bytes = FLOPS_PER_IMAGE * B, HBM_BYTES_PER_IMAGE * B + 8 * PARAM_BYTES

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cpu = cpu_reference_run(B, 2000, 20, budget_s=12.0)
        cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}

    line = {
        "metric": METRIC, "value": world * B * K / (ms_total * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
        "warmup": W, "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"examples/mnist Net DDP train step (BASELINE configs[{1 if world == 1 else 2}]), "
                               f"per-rank batch {B}, synthetic 1x28x28", "global_batch": world * B,
                   "parallelism": f"dp{world}", "l2": f"inputs cycle through {nring} distinct device-resident batches "
                                                      f"({nring * rec / 1e6:.0f} MB > 126 MB L2)",
                   "cuda_graph": True, "lr": 0.01, "momentum": 0.5},
        "e2e": {"value": world * B * K / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": rec,
                "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e / K,
                "how": "StepPipeline.submit(pinned record): H2D on a copy stream into the idle one of two static buffers, graph replay, the loss stored by the last kernel straight into pinned host memory (zero-copy D2H) -- every step"},
        "gpu_launches": kernels_per_step * K,
        "kernels_per_step": kernels_per_step,
        "clocks": clocks, "clocks_e2e": clocks_e2e,
        "roofline": roof,
        "step_roofline": {"flops_per_step": step_flops, "hbm_bytes_per_step": step_bytes,
                          "tflops": step_flops / (step_us * 1e-6) / 1e12,
                          "hbm_gbs": step_bytes / (step_us * 1e-6) / 1e9,
                          "frac_of_bf16_peak": step_flops / (step_us * 1e-6) / 1e12 / peaks["bf16_tflops"],
                          "frac_of_hbm_peak": step_bytes / (step_us * 1e-6) / 1e9 / peaks["hbm_gbs"]},
        "kernel_us": {k: round(v, 3) for k, v in sorted(kern_us.items(), key=lambda kv: -kv[1])},
        "kernel_us_sum": round(sum(kern_us.values()), 3),
        "ddp_exchange": (None if world == 1 else ("fused peer-memory all-reduce+SGD kernel (NVLink P2P, symmetric memory)"
                                                  if model.fused else "NCCL all_reduce + SGD kernel")),
        "grad_allreduce": busbw,
        "cpu_baseline": cpu,
        "final_loss": loss_final,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        net.release_graphs()
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
