"""Probe the tensor-core conv2 data gradient with one-hot gradients (development aid)."""
import os
import sys

import torch

sys.path.insert(0, ".")
from pytorch_operator_b200 import _cabi  # noqa: E402


def run(L, dp2, p2, i2, p1, w, B, mode):
    os.environ["B200MNIST_CONV2_DGRAD_TC"] = mode
    dev = dp2.device
    gw = torch.zeros(50, 20, 5, 5, device=dev)
    gb = torch.zeros(50, device=dev)
    dp1 = torch.full((B, 20, 12, 12), 7.0, device=dev)
    scratch = torch.empty(L.b200mnist_conv2_bwd_scratch_bytes(B), dtype=torch.uint8, device=dev)
    _cabi.check(L.b200mnist_conv2_bwd(dp2.data_ptr(), p2.data_ptr(), i2.data_ptr(), p1.data_ptr(), w.data_ptr(), B,
                                      gw.data_ptr(), gb.data_ptr(), dp1.data_ptr(), scratch.data_ptr(),
                                      torch.cuda.current_stream().cuda_stream), mode)
    torch.cuda.synchronize()
    return dp1.cpu()


def main():
    L = _cabi.lib()
    dev = torch.device("cuda", 0)
    B = 1
    torch.manual_seed(0)
    w = torch.arange(50 * 500, dtype=torch.float32, device=dev).reshape(50, 20, 5, 5) / 1000.0
    p1 = torch.zeros(B, 20, 12, 12, device=dev)
    p2 = torch.ones(B, 50, 4, 4, device=dev)
    i2 = torch.zeros(B, 50, 4, 4, dtype=torch.uint8, device=dev)
    for (oc, p) in [(0, 0), (1, 0), (5, 0), (0, 1), (0, 5), (49, 15)]:
        dp2 = torch.zeros(B, 50, 4, 4, device=dev)
        dp2.view(B, 50, 16)[0, oc, p] = 1.0
        a = run(L, dp2, p2, i2, p1, w, B, "1")
        b = run(L, dp2, p2, i2, p1, w, B, "0")
        d = (a - b).abs()
        print(f"one-hot oc={oc} p={p}: max|tc - simt| = {d.max():.4g}; nonzero tc {int((a != 0).sum())} simt {int((b != 0).sum())}")
        if d.max() > 1e-4:
            nz = (a != 0).nonzero()[:8].tolist()
            print("  tc nonzeros (first 8):", [(i, round(float(a[tuple(i)]), 3)) for i in nz])
            nz = (b != 0).nonzero()[:8].tolist()
            print("  simt nonzeros (first 8):", [(i, round(float(b[tuple(i)]), 3)) for i in nz])


if __name__ == "__main__":
    main()
