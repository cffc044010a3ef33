"""Parity of the CUDA path (through the C ABI of libb200mnist.so) with the CPU oracle and with
the fixtures produced by the unmodified reference.  Run on the B200 box: pytest -m gpu.

Tolerance (north_star: "1e-5 relative on per-step loss and gradients"): |a-b| <= 1e-5*|b| +
1e-5*max|b| per tensor (tests/helpers.py).  SGD and the smoke-dist square are bit-exact."""
import numpy as np
import pytest
import torch

from oracle import c_oracle, mnist_oracle as mo
from tests.helpers import (NAMES, RTOL, assert_close, golden_batch, golden_tensor_check, kink_free_batch, load_golden,
                           tie_free_batch)

pytestmark = pytest.mark.gpu

if torch.cuda.is_available():
    import pytorch_operator_b200 as pkg
    from pytorch_operator_b200 import _cabi
DEV = "cuda:0"


def _net(seed=1):
    torch.manual_seed(seed)
    return pkg.Net().to(DEV)


def _grads(net):
    return {k: v.detach().cpu().numpy() for k, v in net.named_views(net.flat_grads).items()}


def _params(net):
    return {k: v.detach().cpu().numpy() for k, v in net.named_views(net.flat_params).items()}


def _stream():
    return torch.cuda.current_stream().cuda_stream


def test_library_loaded_and_no_cpu_fallback():
    assert _cabi.lib().b200mnist_version() >= 100
    cpu_net = pkg.Net()
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        cpu_net(torch.zeros(1, 1, 28, 28))


@pytest.mark.parametrize("B", [1, 5, 64])
def test_forward_layers_match_c_oracle(B):
    L = _cabi.lib()
    params = {k: v.numpy() for k, v in mo.init_params(1).items()}
    x, _ = mo.synthetic_batch(B, 99 + B)
    act = c_oracle.forward(params, x.numpy())
    d = {k: torch.from_numpy(np.ascontiguousarray(v)).to(DEV) for k, v in params.items()}
    xd = x.to(DEV)
    p1 = torch.empty(B, 20, 12, 12, device=DEV)
    i1 = torch.empty(B, 20, 12, 12, dtype=torch.uint8, device=DEV)
    _cabi.check(L.b200mnist_conv1_relu_pool_fwd(xd.data_ptr(), d["conv1.weight"].data_ptr(), d["conv1.bias"].data_ptr(),
                                                p1.data_ptr(), i1.data_ptr(), B, _stream()), "conv1")
    assert_close(p1.cpu().numpy(), act["p1"], what="p1")
    live = act["p1"] > 1e-4  # arg-max is only defined (and only used) where the ReLU output is positive
    assert np.array_equal(i1.cpu().numpy()[live], act["i1"][live])
    p1o = torch.from_numpy(act["p1"]).to(DEV)
    p2 = torch.empty(B, 50, 4, 4, device=DEV)
    i2 = torch.empty(B, 50, 4, 4, dtype=torch.uint8, device=DEV)
    _cabi.check(L.b200mnist_conv2_relu_pool_fwd(p1o.data_ptr(), d["conv2.weight"].data_ptr(), d["conv2.bias"].data_ptr(),
                                                p2.data_ptr(), i2.data_ptr(), B, _stream()), "conv2")
    assert_close(p2.cpu().numpy(), act["p2"], what="p2")
    live = act["p2"] > 1e-4
    assert np.array_equal(i2.cpu().numpy()[live], act["i2"][live])
    p2o = torch.from_numpy(act["p2"]).to(DEV)
    h = torch.empty(B, 500, device=DEV)
    sc = torch.zeros(int(L.b200mnist_fc1_scratch_bytes(B)), dtype=torch.uint8, device=DEV)
    _cabi.check(L.b200mnist_fc1_relu_fwd(p2o.data_ptr(), d["fc1.weight"].data_ptr(), d["fc1.bias"].data_ptr(),
                                         h.data_ptr(), B, sc.data_ptr(), _stream()), "fc1")
    assert_close(h.cpu().numpy(), act["h"], what="h")


TC_CONV_MIN_BATCH = 128   # mirrors kTcConvMinBatch in csrc/b200mnist.cu


def _gpu_pool_indices(params, x):
    """The arg-max maps the CUDA forward kernels produce for this batch (same kernels, same
    summation order as inside the fused step)."""
    L, B = _cabi.lib(), x.shape[0]
    d = {k: v.to(DEV) for k, v in params.items()}
    xd = x.to(DEV)
    p1 = torch.empty(B, 20, 12, 12, device=DEV)
    i1 = torch.empty(B, 20, 12, 12, dtype=torch.uint8, device=DEV)
    p2 = torch.empty(B, 50, 4, 4, device=DEV)
    i2 = torch.empty(B, 50, 4, 4, dtype=torch.uint8, device=DEV)
    _cabi.check(L.b200mnist_conv1_relu_pool_fwd(xd.data_ptr(), d["conv1.weight"].data_ptr(), d["conv1.bias"].data_ptr(),
                                                p1.data_ptr(), i1.data_ptr(), B, _stream()), "conv1")
    if B >= TC_CONV_MIN_BATCH:   # the fused step runs conv2 on the tensor cores from here on (kTcConvMinBatch)
        sc = torch.empty(L.b200mnist_conv2_fwd_tc_scratch_bytes(), dtype=torch.uint8, device=DEV)
        _cabi.check(L.b200mnist_conv2_relu_pool_fwd_tc(p1.data_ptr(), d["conv2.weight"].data_ptr(), d["conv2.bias"].data_ptr(),
                                                       p2.data_ptr(), i2.data_ptr(), B, sc.data_ptr(), _stream()), "conv2 tc")
    else:
        _cabi.check(L.b200mnist_conv2_relu_pool_fwd(p1.data_ptr(), d["conv2.weight"].data_ptr(), d["conv2.bias"].data_ptr(),
                                                    p2.data_ptr(), i2.data_ptr(), B, _stream()), "conv2")
    return i1.cpu(), i2.cpu()


# 200, 300: conv2 on tcgen05, fc1 on the SIMT GEMMs; 400: fc1 on tcgen05 as well
@pytest.mark.parametrize("B", [1, 2, 5, 63, 64, 65, 200, 300, 400])
def test_forward_backward_matches_oracles(B):
    """Loss, log-probs and all eight gradients of the fused forward+backward.
    (a) On ARBITRARY batches (near-ties in max-pool windows included) against the C oracle
        with the arg-max maps teacher-forced to the CUDA kernels' choice, after checking that
        choice is a valid maximum to fp32 rounding (mnist_oracle.pool_choice_slack).
    (b) On a near-tie-free batch against the torch-CPU oracle, nothing forced."""
    params = mo.init_params(1)
    pn = {k: v.numpy() for k, v in params.items()}
    net = _net()
    # (a)  (from TC_CONV_MIN_BATCH on: a batch whose conv2 ReLU masks are out of reach of the 3xTF32 error)
    if B < TC_CONV_MIN_BATCH:
        x, t = mo.synthetic_batch(B, 500 + B)
    else:   # fc1 joins the tensor-core path at 384 (kTcMinBatch): keep its ReLU masks out of reach as well
        x, t = kink_free_batch(params, B, 500 + B, min_fc1=2e-6 if B >= 384 else 0.0)
    logp = torch.empty(B, 10, device=DEV)
    loss = float(net.forward_backward(x.to(DEV), t.to(DEV), logp_out=logp))
    got = _grads(net)
    i1, i2 = _gpu_pool_indices(params, x)
    assert mo.pool_choice_slack(params, x, i1, i2) <= 2e-6
    lc, logp_c, g_c = c_oracle.loss_and_grads(pn, x.numpy(), t.numpy(), idx_override=(i1.numpy(), i2.numpy()))
    assert abs(loss - lc) <= RTOL * abs(lc) * 2
    assert_close(logp.cpu().numpy(), logp_c, what="logp")
    for n in NAMES:
        assert_close(got[n], g_c[n], what=f"grad {n} (C oracle, B={B})")
    # pad elements of the bucket stay zero
    assert float(net.flat_grads.double().abs().sum()) == pytest.approx(
        sum(float(np.abs(v.astype(np.float64)).sum()) for v in got.values()), rel=1e-9)
    # (b) -- fp32 SIMT path only: the 3xTF32 conv2 (B >= 256) differs from the fp32 one by ~1e-6 relative, and no batch
    # of that size is free of max-pool near-ties at that scale
    if B >= TC_CONV_MIN_BATCH:
        return
    x, t = tie_free_batch(params, B, 9000 + 10 * B, min_gap=5e-7 if B <= 65 else 2e-7)
    loss = float(net.forward_backward(x.to(DEV), t.to(DEV), logp_out=logp))
    lo, logp_o, g_o = mo.loss_and_grads(params, x, t)
    assert abs(loss - float(lo)) <= RTOL * abs(float(lo)) * 2
    assert_close(logp.cpu().numpy(), logp_o.numpy(), what="logp")
    got = _grads(net)
    for n in NAMES:
        assert_close(got[n], g_o[n].numpy(), what=f"grad {n} (torch oracle, B={B})")


def test_step0_gradients_match_reference_fixture():
    gold = load_golden("train_ws1_b64")
    net = _net()
    x, t = golden_batch(gold, 64, 0, 0)
    loss = net.forward_backward(x.to(DEV), t.to(DEV))
    assert abs(float(loss) - float(gold["step0/loss_rank0"][0])) <= RTOL * 2.4
    for n, v in _grads(net).items():
        golden_tensor_check(gold, "step0/grad", n, v)


@pytest.mark.parametrize("case,graph", [("train_ws1_b64", False), ("train_ws1_b64", True), ("train_ws1_b5", False)])
def test_free_running_training_matches_reference_fixture(case, graph):
    """3 (or 1) full steps -- fused forward/backward/SGD, eager and CUDA-graph replay -- against the
    losses, gradients and updated parameters the reference's train() produced."""
    gold = load_golden(case)
    B, nsteps = int(gold["meta/batch_size"][0]), int(gold["meta/nsteps"][0])
    net = _net()
    opt = pkg.SGD(net.parameters(), lr=0.01, momentum=0.5)
    g = None
    for s in range(nsteps):
        x, t = golden_batch(gold, B, s, 0)
        if graph and s > 0:
            g = g or pkg.StepGraph(net, opt, B)
            g.x.copy_(x)
            g.target.copy_(t)
            loss = g.replay()
        else:
            loss = net.train_step(x.to(DEV), t.to(DEV), opt)
        # free-running: step s sees parameters that already carry fp32 noise from earlier steps
        assert abs(float(loss) - float(gold[f"step{s}/loss_rank0"][0])) <= 4 * RTOL * 2.4
        for n, v in _grads(net).items():
            golden_tensor_check(gold, f"step{s}/grad", n, v, rtol=4 * RTOL)
        for n, v in _params(net).items():
            golden_tensor_check(gold, f"step{s}/param", n, v, rtol=RTOL)


def test_reference_shaped_loop_through_autograd_bridge():
    """output = model(data); F.nll_loss(output, target).backward(); optimizer.step() -- the loop
    body of the reference's train() verbatim -- gives the same gradients as the fused call."""
    import torch.nn.functional as F
    gold = load_golden("train_ws1_b64")
    net = _net()
    opt = pkg.SGD(net.parameters(), lr=0.01, momentum=0.5)
    for s in range(2):
        x, t = golden_batch(gold, 64, s, 0)
        data, target = x.to(DEV), t.to(DEV)
        opt.zero_grad()
        output = net(data)
        loss = F.nll_loss(output, target)
        loss.backward()
        for n, p in net.named_parameters():
            golden_tensor_check(gold, f"step{s}/grad", n, p.grad.cpu().numpy(), rtol=4 * RTOL)
        opt.step()
        assert abs(loss.item() - float(gold[f"step{s}/loss_rank0"][0])) <= 4 * RTOL * 2.4
        for n, v in _params(net).items():
            golden_tensor_check(gold, f"step{s}/param", n, v)


def test_sgd_bit_exact_vs_torch_and_c_oracle():
    n = 431296
    g = torch.Generator().manual_seed(3)
    p0 = torch.randn(n, generator=g)
    p_t = p0.clone().requires_grad_(True)
    opt = torch.optim.SGD([p_t], lr=0.01, momentum=0.5)
    p_d, m_d = p0.to(DEV), torch.zeros(n, device=DEV)
    p_c, m_c = p0.numpy().copy(), np.zeros(n, np.float32)
    L = _cabi.lib()
    for step in range(3):
        grad = torch.randn(n, generator=g)
        p_t.grad = (grad * 0.25).clone()  # DDP mean over 4 ranks: exact power-of-two scale
        opt.step()
        gd = grad.to(DEV)
        _cabi.check(L.b200mnist_sgd_momentum(p_d.data_ptr(), gd.data_ptr(), m_d.data_ptr(), n, 0.01, 0.5, 0.25,
                                             1 if step == 0 else 0, _stream()), "sgd")
        c_oracle.sgd_momentum_(p_c, grad.numpy(), m_c, 0.01, 0.5, 0.25, step == 0)
        assert torch.equal(p_d.cpu(), p_t.detach()), f"step {step} vs torch.optim.SGD"
        assert np.array_equal(p_d.cpu().numpy(), p_c), f"step {step} vs C oracle"
    # ragged length / empty input
    for m in (0, 1, 3, 1023, 1025):
        a, b, c = torch.ones(max(m, 4), device=DEV), torch.ones(max(m, 4), device=DEV), torch.zeros(max(m, 4), device=DEV)
        _cabi.check(L.b200mnist_sgd_momentum(a.data_ptr(), b.data_ptr(), c.data_ptr(), m, 0.5, 0.0, 1.0, 1, _stream()), "sgd")
        assert torch.equal(a.cpu()[:m], torch.full((m,), 0.5)) and torch.equal(a.cpu()[m:], torch.ones(max(m, 4) - m))


def test_eval_matches_reference_fixture():
    gold = load_golden("eval_b1000")
    net = _net()
    net.eval()
    totals = torch.zeros(2, dtype=torch.float64, device=DEV)
    for i, seed in enumerate((4321, 4322)):
        x, t = mo.synthetic_batch(1000, seed)
        if i == 0:
            with torch.no_grad():
                logp = net(x.to(DEV))
            assert_close(logp.cpu().numpy(), gold["logp_batch0"], what="logp")
            nll_o, correct_o, _ = mo.evaluate(mo.init_params(1), x, t)
            st = net.evaluate(x.to(DEV), t.to(DEV)).cpu()
            assert abs(float(st[0]) - nll_o) <= RTOL * nll_o * 2 and int(st[1]) == correct_o
        totals += net.evaluate(x.to(DEV), t.to(DEV)).double()
    assert float(totals[1]) / 2000.0 == float(gold["accuracy"][0])


def test_square_bit_exact_and_smoke_fixture():
    from pytorch_operator_b200.dist_sendrecv import square
    gold = load_golden("smoke_sendrecv_ws3")
    inp = torch.from_numpy(gold["inp"])
    out = square(inp.to(DEV)).cpu()
    assert np.array_equal(out.numpy(), gold["result_from_worker1"])
    g = torch.Generator().manual_seed(11)
    for n in (1, 255, 257, 100003):
        v = torch.randn(n, generator=g) * 1e3
        assert torch.equal(square(v.to(DEV)).cpu(), torch.mul(v, v))
    assert square(torch.empty(0, device=DEV)).numel() == 0


def test_full_size_batch_properties():
    """BASELINE config 5 size (B=4096, where fc1 runs on the tcgen05 tensor-core path): results are checked
    through size-independent properties -- bitwise repeatability, the batch gradient is the mean of the
    chunk gradients (linearity of the mean loss; chunks of 1024 so every chunk takes the same kernels as
    the full batch), the loss is the mean of the chunk losses -- and the forward of the big batch is compared
    with the same images run as a small batch through the fp32 SIMT path.
    (Gradients are NOT compared across the two fc1 paths: an fc1 output within rounding of zero can land on
    either side of the ReLU kink, which moves individual gradient entries by O(1/B) -- the same kind of
    legitimate discontinuity as the max-pool near-ties.)"""
    B, C = 4096, 1024
    net = _net()
    x, t = mo.synthetic_batch(B, 777)
    xd, td = x.to(DEV), t.to(DEV)
    lp_big = torch.empty(B, 10, device=DEV)
    loss = float(net.forward_backward(xd, td, logp_out=lp_big))
    g_full = net.flat_grads.clone()
    loss2 = float(net.forward_backward(xd, td))
    assert loss == loss2 and torch.equal(g_full, net.flat_grads), "not bitwise repeatable"
    acc = torch.zeros_like(g_full, dtype=torch.float64)
    lsum = 0.0
    for c in range(0, B, C):
        lsum += float(net.forward_backward(xd[c:c + C], td[c:c + C]))
        acc += net.flat_grads.double()
    assert abs(loss - lsum / (B // C)) <= RTOL * abs(loss)
    assert_close(g_full.cpu().numpy(), (acc / (B // C)).cpu().numpy(), rtol=2 * RTOL, what="batch-additivity")
    # tensor-core path (B=4096) vs fp32 SIMT path (B=512) on the same images, and vs the CPU oracle
    lp_small = torch.empty(512, 10, device=DEV)
    net.forward_backward(xd[:512], td[:512], logp_out=lp_small)
    assert_close(lp_big[:512].cpu().numpy(), lp_small.cpu().numpy(), what="logp: tcgen05 path vs SIMT path")
    lp_oracle = mo.forward(mo.init_params(1), x[:512])
    assert_close(lp_big[:512].cpu().numpy(), lp_oracle.numpy(), what="logp: tcgen05 path vs torch oracle")


def _saved_activations(net, B):
    """p1, i1, p2, i2, h as the fused step left them in the workspace (b200mnist_workspace_offset)."""
    L, ws = _cabi.lib(), net.workspace(B)
    def region(which, nbytes):
        off = int(L.b200mnist_workspace_offset(B, which))
        assert off >= 0
        return ws[off:off + nbytes]
    p1 = region(0, B * 2880 * 4).view(torch.float32).view(B, 20, 12, 12)
    i1 = region(1, B * 2880).view(B, 20, 12, 12)
    p2 = region(2, B * 800 * 4).view(torch.float32).view(B, 50, 4, 4)
    i2 = region(3, B * 800).view(B, 50, 4, 4)
    h = region(4, B * 500 * 4).view(torch.float32).view(B, 500)
    return [v.cpu().numpy().copy() for v in (p1, i1, p2, i2, h)]


@pytest.mark.parametrize("B,seed", [(4096, 777), (4096, 31337), (1000, 5), (400, 999), (200, 12345), (128, 3)])
def test_large_batch_gradients_match_c_oracle_on_arbitrary_batches(B, seed):
    """BASELINE configs[4] (B=4096: conv2 forward / data gradient / weight gradient and all three fc1 GEMMs on the
    tcgen05 3xTF32 kernels) and the smaller batches where the tensor-core kernels switch on (conv2 from 128, fc1 from
    384) against the C oracle on ARBITRARY batches -- no search for a convenient seed.
    The backward takes discrete decisions (max-pool arg-max, three ReLU masks) that an implementation accurate to
    ~1e-6 can legitimately take differently where the deciding value is within rounding of a tie / of zero.  So the
    oracle's backward is teacher-forced with the CUDA run's arg-max maps and masks -- after checking that every
    forced arg-max is a maximum to 2e-6 (float64) and that every forced ReLU mask belongs to an activation within
    5e-6 of the kink -- and then loss, log-probs and all eight gradients must agree to the 1e-5 of north_star.
    The oracle runs in chunks of 256 images on a thread pool (the mean loss is linear in the chunks)."""
    from concurrent.futures import ThreadPoolExecutor
    params = mo.init_params(1)
    pn = {k: v.numpy() for k, v in params.items()}
    net = _net()
    x, t = mo.synthetic_batch(B, seed)
    logp = torch.empty(B, 10, device=DEV)
    loss = float(net.forward_backward(x.to(DEV), t.to(DEV), logp_out=logp))
    got = _grads(net)
    p1, i1, p2, i2, h = _saved_activations(net, B)
    C = 256
    chunks = [(c, min(B, c + C)) for c in range(0, B, C)]
    xn, tn = x.numpy(), t.numpy()

    def one(ch):
        a, b = ch
        slack = mo.pool_choice_slack(params, x[a:b], torch.from_numpy(i1[a:b]), torch.from_numpy(i2[a:b]))
        return (b - a, slack) + c_oracle.loss_and_grads(pn, xn[a:b], tn[a:b], idx_override=(i1[a:b], i2[a:b]),
                                                        mask_override=(p1[a:b], p2[a:b], h[a:b]))

    with ThreadPoolExecutor(max_workers=8) as pool:
        res = list(pool.map(one, chunks))
    assert max(r[1] for r in res) <= 2e-6, "a forced arg-max is not a maximum"
    nforced = {k: sum(r[5][k][0] for r in res) for k in ("p1", "p2", "h")}
    worst = {k: max(r[5][k][1] for r in res) for k in ("p1", "p2", "h")}
    assert max(worst.values()) <= 5e-6, (nforced, worst)
    assert sum(nforced.values()) <= 1e-4 * B * (2880 + 800 + 500), nforced   # a handful, not a systematic difference
    lc = sum(r[0] * r[2] for r in res) / B
    assert abs(loss - lc) <= RTOL * abs(lc) * 2
    assert_close(logp.cpu().numpy(), np.concatenate([r[3] for r in res]), what="logp")
    for n in NAMES:
        ref = sum(r[4][n].astype(np.float64) * (r[0] / B) for r in res)
        assert_close(got[n], ref, what=f"grad {n} (C oracle, forced choices, B={B}, forced masks {nforced})")


def test_argument_errors_and_bad_target():
    L = _cabi.lib()
    net = _net()
    x, t = mo.synthetic_batch(4, 5)
    xd, td = x.to(DEV), t.to(DEV)
    ws = net.workspace(4)
    rc = L.b200mnist_forward_backward(net.flat_params.data_ptr(), net.flat_grads.data_ptr(), xd.data_ptr(),
                                      td.data_ptr(), 8, ws.data_ptr(), ws.numel(), net._loss.data_ptr(), 0, _stream())
    assert rc == -10004 and "workspace" in _cabi.last_error()
    rc = L.b200mnist_forward_backward(0, net.flat_grads.data_ptr(), xd.data_ptr(), td.data_ptr(), 4, ws.data_ptr(),
                                      ws.numel(), net._loss.data_ptr(), 0, _stream())
    assert rc == -10001
    rc = L.b200mnist_forward_backward(net.flat_params.data_ptr(), net.flat_grads.data_ptr(), xd.data_ptr(),
                                      td.data_ptr(), 0, ws.data_ptr(), ws.numel(), net._loss.data_ptr(), 0, _stream())
    assert rc == -10002
    with pytest.raises(ValueError):
        net.forward_backward(xd.double(), td)
    td_bad = td.clone()
    td_bad[2] = 10  # torch raises "Target 10 is out of bounds"; the kernel poisons the loss instead
    assert np.isnan(float(net.forward_backward(xd, td_bad)))
    assert np.isfinite(float(net.forward_backward(xd, td)))


def test_state_dict_roundtrip_and_device_move():
    net = _net()
    sd = {k: v.detach().cpu().clone() for k, v in net.state_dict().items()}
    assert list(sd) == NAMES
    ref = mo.init_params(1)
    assert all(torch.equal(sd[k], ref[k]) for k in NAMES)
    other = _net(seed=2)
    other.load_named(sd)
    assert torch.equal(other.flat_params, net.flat_params)
    for n, p in net.named_parameters():  # views alias the bucket after .to()
        assert p.data_ptr() >= net.flat_params.data_ptr() and p.grad.data_ptr() >= net.flat_grads.data_ptr()
