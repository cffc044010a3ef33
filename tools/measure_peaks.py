#!/usr/bin/env python
"""Measure the two compute rooflines MEASURED_PEAKS.json does not hold, the way that file was made (best of 10,
CUDA events, after warm-up), and write profiles/r02_peaks_fp32_tf32.json:

  fp32_ffma_tflops : libb200mnist's register-resident packed-FMA loop (fma.rn.f32x2; b200mnist_fma_peak) on
                     148 x 8 CTAs of 256 threads -- the ceiling of every fp32 SIMT kernel in this repo;
  tf32_tflops      : torch.matmul fp32 8192^3 with allow_tf32 (cuBLAS TF32 tensor-core GEMM) -- the ceiling of the
                     tcgen05 kind::tf32 kernels; the 3xTF32 product does 3 (or 4) MMAs per algorithmic one, so those
                     kernels are budgeted against tf32_tflops / 3;
  fp32_matmul_tflops : the same matmul without TF32 (cuBLAS SGEMM), for reference.

    python tools/measure_peaks.py [out.json]
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch

import pytorch_operator_b200 as pkg


def best_of(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    best = float("inf")
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e-3)
    return best


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_peaks_fp32_tf32.json")
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    lib = pkg._cabi.lib()
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    blocks, iters = sms * 8, 1 << 15
    out = torch.empty(blocks * 256, dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    flops = [0]

    def fma():
        flops[0] = int(lib.b200mnist_fma_peak(out.data_ptr(), blocks, iters, stream))
        assert flops[0] > 0, pkg._cabi.last_error()

    t = best_of(fma)
    res = {"fp32_ffma_tflops": flops[0] / t / 1e12, "fp32_ffma_how": f"fma.rn.f32x2 loop, {blocks} CTAs x 256 threads x "
           f"{iters} passes x 16 packed FMAs = {flops[0]:.3e} FLOP in {t * 1e3:.3f} ms, best of 10"}
    n = 8192
    a = torch.randn(n, n, device=dev)
    b = torch.randn(n, n, device=dev)
    for key, tf32 in (("tf32_tflops", True), ("fp32_matmul_tflops", False)):
        torch.backends.cuda.matmul.allow_tf32 = tf32
        t = best_of(lambda: torch.matmul(a, b))
        res[key] = 2.0 * n ** 3 / t / 1e12
    torch.backends.cuda.matmul.allow_tf32 = False
    res.update({"tf32_how": "torch.matmul fp32 8192^3, allow_tf32=True, best of 10 (cuBLAS)",
                "gpu_name": torch.cuda.get_device_name(dev), "sms": sms, "torch": torch.__version__,
                "when": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime())})
    with open(out_path, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
