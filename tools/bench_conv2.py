"""Time the two conv2 forward entry points (FFMA2 direct convolution vs tcgen05 shifted-window GEMM) stand-alone.

    python tools/bench_conv2.py [B ...]
"""
import sys

import torch

sys.path.insert(0, ".")
from pytorch_operator_b200 import _cabi  # noqa: E402


def main():
    L = _cabi.lib()
    dev = torch.device("cuda", 0)
    sizes = [int(a) for a in sys.argv[1:]] or [64, 256, 1024, 4096]
    w = torch.randn(50, 20, 5, 5, device=dev) * 0.05
    b = torch.randn(50, device=dev) * 0.05
    scratch = torch.empty(L.b200mnist_conv2_fwd_tc_scratch_bytes(), dtype=torch.uint8, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    for B in sizes:
        p1 = torch.relu(torch.randn(B, 20, 12, 12, device=dev))
        out = torch.empty(B, 50, 4, 4, device=dev)
        idx = torch.empty(B, 50, 4, 4, dtype=torch.uint8, device=dev)
        res = {}
        for name in ("simt", "tc"):
            def call():
                if name == "tc":
                    rc = L.b200mnist_conv2_relu_pool_fwd_tc(p1.data_ptr(), w.data_ptr(), b.data_ptr(), out.data_ptr(),
                                                            idx.data_ptr(), B, scratch.data_ptr(), s)
                else:
                    rc = L.b200mnist_conv2_relu_pool_fwd(p1.data_ptr(), w.data_ptr(), b.data_ptr(), out.data_ptr(),
                                                         idx.data_ptr(), B, s)
                _cabi.check(rc, name)
            for _ in range(5):
                call()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n = 50
            e0.record()
            for _ in range(n):
                call()
            e1.record()
            torch.cuda.synchronize()
            res[name] = e0.elapsed_time(e1) / n * 1e3
            res[name + "_sum"] = float(out.double().sum())
        print(f"B={B}: simt {res['simt']:.1f} us  tc (split + conv) {res['tc']:.1f} us   checksums {res['simt_sum']:.4f} {res['tc_sum']:.4f}",
              flush=True)


if __name__ == "__main__":
    main()
