"""Time the conv2 backward entry point (b200mnist_conv2_bwd) with the tensor-core / SIMT halves toggled.

    python tools/bench_conv2_bwd.py [B ...]
"""
import os
import sys

import torch

sys.path.insert(0, ".")
from pytorch_operator_b200 import _cabi  # noqa: E402


def main():
    L = _cabi.lib()
    dev = torch.device("cuda", 0)
    sizes = [int(a) for a in sys.argv[1:]] or [64, 1024, 4096]
    w = torch.randn(50, 20, 5, 5, device=dev) * 0.05
    b = torch.randn(50, device=dev) * 0.05
    s = torch.cuda.current_stream().cuda_stream
    for B in sizes:
        p1 = torch.relu(torch.randn(B, 20, 12, 12, device=dev))
        p2 = torch.empty(B, 50, 4, 4, device=dev)
        i2 = torch.empty(B, 50, 4, 4, dtype=torch.uint8, device=dev)
        _cabi.check(L.b200mnist_conv2_relu_pool_fwd(p1.data_ptr(), w.data_ptr(), b.data_ptr(), p2.data_ptr(), i2.data_ptr(), B, s), "fwd")
        dp2 = torch.randn(B, 50, 4, 4, device=dev)
        gw, gb = torch.empty(50, 20, 5, 5, device=dev), torch.empty(50, device=dev)
        dp1 = torch.empty(B, 20, 12, 12, device=dev)
        scratch = torch.empty(L.b200mnist_conv2_bwd_scratch_bytes(B), dtype=torch.uint8, device=dev)
        out = []
        for dg, wg in (("0", "0"), ("1", "0"), ("0", "1"), ("1", "1")):
            os.environ["B200MNIST_CONV2_DGRAD_TC"], os.environ["B200MNIST_CONV2_WGRAD_TC"] = dg, wg

            def call():
                _cabi.check(L.b200mnist_conv2_bwd(dp2.data_ptr(), p2.data_ptr(), i2.data_ptr(), p1.data_ptr(), w.data_ptr(), B,
                                                  gw.data_ptr(), gb.data_ptr(), dp1.data_ptr(), scratch.data_ptr(), s), "bwd")
            for _ in range(3):
                call()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n = 20
            e0.record()
            for _ in range(n):
                call()
            e1.record()
            torch.cuda.synchronize()
            out.append(f"dgrad_tc={dg} wgrad_tc={wg}: {e0.elapsed_time(e1) / n * 1e3:.1f} us")
        print(f"B={B}: " + "; ".join(out) + "  (includes the two partial-sum reductions)", flush=True)


if __name__ == "__main__":
    main()
