"""The tcgen05 (5th-gen tensor core) 3xTF32 fc1 GEMM against a float64 reference: fp32-level accuracy
(north_star tolerance 1e-5) despite TF32 inputs, including ragged M (batch not a multiple of 128) and the
N tail (500 = 7 x 64 + 52)."""
import pytest
import torch

from tests.helpers import RTOL, assert_close

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("B", [128, 1, 200, 1000, 4096])
def test_fc1_tensor_core_path_matches_fp64(B):
    from pytorch_operator_b200 import _cabi
    L = _cabi.lib()
    g = torch.Generator().manual_seed(40 + B)
    x = torch.relu(torch.randn(B, 800, generator=g))          # like the pooled conv2 activations: >= 0, many zeros
    w = (torch.rand(500, 800, generator=g) - 0.5) * 0.07      # fc1's default-init scale
    b = (torch.rand(500, generator=g) - 0.5) * 0.07
    xd, wd, bd = x.to(DEV), w.to(DEV), b.to(DEV)
    out = torch.full((B, 500), 7.0, device=DEV)
    sc = torch.zeros(int(L.b200mnist_fc1_scratch_bytes(B)), dtype=torch.uint8, device=DEV)
    _cabi.check(L.b200mnist_fc1_relu_fwd_tc(xd.data_ptr(), wd.data_ptr(), bd.data_ptr(), out.data_ptr(), B, sc.data_ptr(),
                                            torch.cuda.current_stream().cuda_stream), "fc1 tc")
    torch.cuda.synchronize()
    ref = torch.relu(x.double() @ w.double().T + b.double())
    assert_close(out.cpu().numpy(), ref.numpy(), rtol=RTOL, what=f"fc1 tcgen05 B={B}")
    # and against the fp32 SIMT kernel (same tolerance)
    out2 = torch.empty(B, 500, device=DEV)
    if B < 384:  # b200mnist_fc1_relu_fwd itself switches to the tensor-core path from 384 up (kTcMinBatch)
        _cabi.check(L.b200mnist_fc1_relu_fwd(xd.data_ptr(), wd.data_ptr(), bd.data_ptr(), out2.data_ptr(), B,
                                             sc.data_ptr(), torch.cuda.current_stream().cuda_stream), "fc1 simt")
        assert_close(out.cpu().numpy(), out2.cpu().numpy(), rtol=RTOL, what="tcgen05 vs SIMT")


@pytest.mark.parametrize("B", [1024, 1500, 4096])
def test_fc_backward_tensor_core_path_matches_fp64(B):
    """b200mnist_fc_bwd at B >= 1024: fc1 weight gradient (MN-major x MN-major, bias gradient through an injected
    column of ones) and fc1 data gradient (K-major x MN-major) run on tcgen05 3xTF32; fc2 stays on the SIMT kernel."""
    from pytorch_operator_b200 import _cabi
    L = _cabi.lib()
    g = torch.Generator().manual_seed(90 + B)
    dz2 = (torch.randn(B, 10, generator=g) / B)
    dz1 = (torch.randn(B, 500, generator=g) / B) * (torch.rand(B, 500, generator=g) > 0.5)   # ReLU-masked, like the step
    h = torch.relu(torch.randn(B, 500, generator=g))
    p2 = torch.relu(torch.randn(B, 800, generator=g))
    w1 = (torch.rand(500, 800, generator=g) - 0.5) * 0.07
    d = {k: v.to(DEV) for k, v in dict(dz2=dz2, dz1=dz1, h=h, p2=p2, w1=w1).items()}
    out = {"gw2": torch.full((10, 500), 7.0, device=DEV), "gb2": torch.full((10,), 7.0, device=DEV),
           "gw1": torch.full((500, 800), 7.0, device=DEV), "gb1": torch.full((500,), 7.0, device=DEV),
           "dp2": torch.full((B, 800), 7.0, device=DEV)}
    sc = torch.zeros(int(L.b200mnist_fc_bwd_scratch_bytes(B)), dtype=torch.uint8, device=DEV)
    _cabi.check(L.b200mnist_fc_bwd(d["dz2"].data_ptr(), d["dz1"].data_ptr(), d["h"].data_ptr(), d["p2"].data_ptr(),
                                   d["w1"].data_ptr(), B, out["gw2"].data_ptr(), out["gb2"].data_ptr(),
                                   out["gw1"].data_ptr(), out["gb1"].data_ptr(), out["dp2"].data_ptr(), sc.data_ptr(),
                                   torch.cuda.current_stream().cuda_stream), "fc_bwd")
    torch.cuda.synchronize()
    ref = {"gw2": dz2.double().T @ h.double(), "gb2": dz2.double().sum(0), "gw1": dz1.double().T @ p2.double(),
           "gb1": dz1.double().sum(0), "dp2": dz1.double() @ w1.double()}
    for k in ref:
        assert_close(out[k].cpu().numpy(), ref[k].numpy(), rtol=RTOL, what=f"{k} (B={B})")


@pytest.mark.parametrize("B", [1, 2, 63, 64, 300, 1000, 4096])
def test_conv2_implicit_gemm_tensor_core_path_matches_oracle(B):
    """conv2 + ReLU + pool as a tcgen05 shifted-window implicit GEMM against float64, and its arg-max map against
    the fp32 SIMT kernel wherever the window maximum is not a near-tie.  B=300 and 1000 make the persistent CTAs
    walk several image pairs (double-buffered staging / accumulators)."""
    import numpy as np
    import torch.nn.functional as F
    from oracle import mnist_oracle as mo
    from pytorch_operator_b200 import _cabi
    L = _cabi.lib()
    params = mo.init_params(1)
    g = torch.Generator().manual_seed(70 + B)
    p1 = torch.relu(torch.randn(B, 20, 12, 12, generator=g))
    w, b = params["conv2.weight"], params["conv2.bias"]
    d = {"p1": p1.to(DEV), "w": w.to(DEV), "b": b.to(DEV)}
    outs = {}
    scratch = torch.empty(L.b200mnist_conv2_fwd_tc_scratch_bytes(), dtype=torch.uint8, device=DEV)
    for name, fn in (("tc", L.b200mnist_conv2_relu_pool_fwd_tc), ("simt", L.b200mnist_conv2_relu_pool_fwd)):
        out = torch.full((B, 50, 4, 4), 7.0, device=DEV)
        idx = torch.full((B, 50, 4, 4), 9, dtype=torch.uint8, device=DEV)
        extra = (scratch.data_ptr(),) if name == "tc" else ()
        _cabi.check(fn(d["p1"].data_ptr(), d["w"].data_ptr(), d["b"].data_ptr(), out.data_ptr(), idx.data_ptr(), B,
                       *extra, torch.cuda.current_stream().cuda_stream), name)
        torch.cuda.synchronize()
        outs[name] = (out.cpu(), idx.cpu())
    c = F.relu(F.conv2d(p1.to(DEV).double(), w.to(DEV).double(), b.to(DEV).double())).cpu()   # float64 on the GPU
    ref = F.max_pool2d(c, 2, 2)
    assert_close(outs["tc"][0].numpy(), ref.numpy(), rtol=RTOL, what=f"p2 tcgen05 B={B}")
    win = c.unfold(2, 2, 2).unfold(3, 2, 2).reshape(B, 50, 4, 4, 4)
    top2 = win.topk(2, dim=-1).values
    clear = ((top2[..., 0] - top2[..., 1]) > 1e-5) & (top2[..., 0] > 1e-5)
    assert np.array_equal(outs["tc"][1].numpy()[clear.numpy()], win.argmax(-1).numpy()[clear.numpy()])
    assert np.array_equal(outs["tc"][1].numpy()[clear.numpy()], outs["simt"][1].numpy()[clear.numpy()])


@pytest.mark.parametrize("B", [1, 2, 3, 64, 75, 300, 4096])
def test_conv2_backward_tensor_core_paths_match_float64(B, monkeypatch):
    """conv2 backward with the data gradient (GEMM + col2im epilogue) and/or the weight gradient (one MMA per
    hi/lo quartet, fp32 folds every 8 images) on tcgen05, and with the fp32 SIMT kernel: all four combinations
    against float64 autograd formulas on the same arg-max map.  Odd B leaves half an image pair empty, B=75/300
    make the persistent CTAs walk several tiles (double-buffered operands, patches, accumulators, drains); B=4096
    (BASELINE configs[4]) runs the weight gradient's 29 slices x ~141 images with a TMEM -> shared-memory fold every
    8 images and ~55 data-gradient tiles per CTA."""
    import torch.nn.functional as F
    from oracle import mnist_oracle as mo
    from pytorch_operator_b200 import _cabi
    L = _cabi.lib()
    params = mo.init_params(3)
    g = torch.Generator().manual_seed(900 + B)
    p1 = torch.relu(torch.randn(B, 20, 12, 12, generator=g)).to(DEV)
    w, b = params["conv2.weight"].to(DEV), params["conv2.bias"].to(DEV)
    dp2 = torch.randn(B, 50, 4, 4, generator=g).to(DEV)
    s = torch.cuda.current_stream().cuda_stream
    p2 = torch.empty(B, 50, 4, 4, device=DEV)
    i2 = torch.empty(B, 50, 4, 4, dtype=torch.uint8, device=DEV)
    _cabi.check(L.b200mnist_conv2_relu_pool_fwd(p1.data_ptr(), w.data_ptr(), b.data_ptr(), p2.data_ptr(), i2.data_ptr(), B, s), "fwd")
    # float64 reference on the kernel's own arg-max map: dconv2 = unpool(dp2 * [p2 > 0])
    dc2 = torch.zeros(B, 50, 8, 8, dtype=torch.float64, device=DEV)
    py, px = torch.meshgrid(torch.arange(4, device=DEV), torch.arange(4, device=DEV), indexing="ij")
    oy = (2 * py)[None, None] + (i2.long() >> 1)
    ox = (2 * px)[None, None] + (i2.long() & 1)
    bi = torch.arange(B, device=DEV)[:, None, None, None].expand_as(oy)
    ci = torch.arange(50, device=DEV)[None, :, None, None].expand_as(oy)
    dc2[bi, ci, oy, ox] = torch.where(p2 > 0, dp2, torch.zeros_like(dp2)).double()
    ref_dp1 = F.conv_transpose2d(dc2, w.double())
    ref_gw = torch.nn.grad.conv2d_weight(p1.double(), w.shape, dc2)
    ref_gb = dc2.sum((0, 2, 3))
    scratch = torch.empty(L.b200mnist_conv2_bwd_scratch_bytes(B), dtype=torch.uint8, device=DEV)
    counts = {}
    for dg, wg in (("0", "0"), ("1", "0"), ("0", "1"), ("1", "1")):
        monkeypatch.setenv("B200MNIST_CONV2_DGRAD_TC", dg)
        monkeypatch.setenv("B200MNIST_CONV2_WGRAD_TC", wg)
        gw = torch.full((50, 20, 5, 5), 7.0, device=DEV)
        gb = torch.full((50,), 7.0, device=DEV)
        dp1 = torch.full((B, 20, 12, 12), 7.0, device=DEV)
        n0 = L.b200mnist_launch_count()
        _cabi.check(L.b200mnist_conv2_bwd(dp2.data_ptr(), p2.data_ptr(), i2.data_ptr(), p1.data_ptr(), w.data_ptr(), B,
                                          gw.data_ptr(), gb.data_ptr(), dp1.data_ptr(), scratch.data_ptr(), s), dg + wg)
        torch.cuda.synchronize()
        counts[dg + wg] = L.b200mnist_launch_count() - n0
        what = f"dgrad tc={dg} wgrad tc={wg} B={B}"
        assert_close(dp1.cpu().numpy(), ref_dp1.cpu().numpy(), rtol=RTOL, what="dp1 " + what)
        assert_close(gw.cpu().numpy(), ref_gw.cpu().numpy(), rtol=RTOL, what="gw " + what)
        assert_close(gb.cpu().numpy(), ref_gb.cpu().numpy(), rtol=RTOL, what="gb " + what)
    # merged SIMT launch -> (weight split + dgrad GEMM + SIMT wgrad) / (tcgen05 wgrad + SIMT dgrad) / (split + both GEMMs)
    assert [counts[k] - counts["00"] for k in ("10", "01", "11")] == [2, 1, 2]
