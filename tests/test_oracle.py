"""Pins the two CPU restatements (oracle/mnist_oracle.py on torch ops, oracle/lenet_ref.c in
plain C) against the fixtures the UNMODIFIED reference produced (oracle/gen_golden.py)."""
import numpy as np
import pytest
import torch

from oracle import c_oracle, mnist_oracle as mo
from tests.helpers import NAMES, RTOL, assert_close, golden_batch, golden_tensor_check, load_golden


def test_param_count_and_init_matches_reference():
    gold = load_golden("train_ws1_b64")
    params = mo.init_params(1)
    assert mo.N_PARAMS == 431080 == sum(v.numel() for v in params.values())
    for n in NAMES:
        golden_tensor_check(gold, "init", n, params[n].numpy(), rtol=0.0)  # bit-exact


@pytest.mark.parametrize("case", ["train_ws1_b64", "train_ws1_b5", "train_ws2_b64"])
def test_torch_restatement_free_run_matches_reference(case):
    gold = load_golden(case)
    ws, B, nsteps = (int(gold[f"meta/{k}"][0]) for k in ("world_size", "batch_size", "nsteps"))
    params = mo.init_params(1)
    batches = [[golden_batch(gold, B, s, r) for s in range(nsteps)] for r in range(ws)]
    trace, _ = mo.train_steps(params, batches)
    # replay again to capture params after each step
    params = mo.init_params(1)
    bufs = {}
    for s in range(nsteps):
        for r in range(ws):
            assert abs(trace[s]["losses"][r] - float(gold[f"step{s}/loss_rank{r}"][0])) <= RTOL * 2.4
        for n in NAMES:
            golden_tensor_check(gold, f"step{s}/grad", n, trace[s]["grads"][n].numpy())
        mo.sgd_momentum_(params, trace[s]["grads"], bufs)
        for n in NAMES:
            golden_tensor_check(gold, f"step{s}/param", n, params[n].numpy())


@pytest.mark.parametrize("case", ["train_ws1_b64", "train_ws1_b5"])
def test_c_restatement_step0_matches_reference(case):
    gold = load_golden(case)
    B = int(gold["meta/batch_size"][0])
    params = {k: v.numpy() for k, v in mo.init_params(1).items()}
    x, t = golden_batch(gold, B, 0, 0)
    loss, logp, grads = c_oracle.loss_and_grads(params, x.numpy(), t.numpy())
    assert abs(loss - float(gold["step0/loss_rank0"][0])) <= RTOL * 2.4
    for n in NAMES:
        golden_tensor_check(gold, "step0/grad", n, grads[n])


def test_c_and_torch_restatements_agree_ws2_average():
    gold = load_golden("train_ws2_b64")
    params = {k: v.numpy() for k, v in mo.init_params(1).items()}
    gsum = None
    for r in range(2):
        x, t = golden_batch(gold, 64, 0, r)
        loss, _, g = c_oracle.loss_and_grads(params, x.numpy(), t.numpy())
        assert abs(loss - float(gold[f"step0/loss_rank{r}"][0])) <= RTOL * 2.4
        gsum = g if gsum is None else {n: gsum[n] + g[n] for n in NAMES}
    for n in NAMES:
        golden_tensor_check(gold, "step0/grad", n, gsum[n] * np.float32(0.5))


def test_sgd_restatements_bit_exact_vs_torch_optim():
    """lenet_sgd_momentum must round exactly like torch.optim.SGD on CPU (3 steps)."""
    g = torch.Generator().manual_seed(7)
    p0 = torch.randn(10007, generator=g)
    p_t = p0.clone().requires_grad_(True)
    opt = torch.optim.SGD([p_t], lr=0.01, momentum=0.5)
    p_c = p0.numpy().copy()
    buf_c = np.zeros_like(p_c)
    for step in range(3):
        grad = torch.randn(10007, generator=g)
        p_t.grad = grad.clone()
        opt.step()
        c_oracle.sgd_momentum_(p_c, grad.numpy(), buf_c, 0.01, 0.5, 1.0, step == 0)
        assert np.array_equal(p_c, p_t.detach().numpy()), f"step {step}"


def test_eval_matches_reference():
    gold = load_golden("eval_b1000")
    params = mo.init_params(1)
    correct = 0
    for i, seed in enumerate((4321, 4322)):
        x, t = mo.synthetic_batch(1000, seed)
        _, c, logp = mo.evaluate(params, x, t)
        correct += c
        if i == 0:
            assert_close(logp.numpy(), gold["logp_batch0"], what="logp torch")
            act = c_oracle.forward({k: v.numpy() for k, v in params.items()}, x.numpy()[:64])
            assert_close(act["logp"], gold["logp_batch0"][:64], what="logp C")
    assert correct / 2000.0 == float(gold["accuracy"][0])


def test_smoke_sendrecv_bit_exact():
    gold = load_golden("smoke_sendrecv_ws3")
    torch.manual_seed(0)
    inp = torch.randn(2, 2)
    assert np.array_equal(inp.numpy(), gold["inp"])
    for w, res in enumerate(mo.smoke_sendrecv(inp, 3), start=1):
        assert np.array_equal(res.numpy(), gold[f"result_from_worker{w}"])
        assert np.array_equal(c_oracle.square(inp.numpy()), gold[f"result_from_worker{w}"])


def test_kink_and_tie_gaps_select_usable_batches():
    """The batch selectors the GPU parity tests rely on: pool_tie_gap / relu_kink_gap are positive distances, a batch
    returned by tie_free_batch / kink_free_batch honours the bound, and the gaps react to a planted near-tie."""
    from tests.helpers import kink_free_batch, tie_free_batch
    params = mo.init_params(1)
    x, _ = tie_free_batch(params, 4, 100, min_gap=5e-7)
    assert mo.pool_tie_gap(params, x) >= 5e-7
    x, _ = kink_free_batch(params, 8, 200, min_conv2=5e-6, min_fc1=2e-6)
    g2, g1 = mo.relu_kink_gap(params, x)
    assert g2 >= 5e-6 and g1 >= 2e-6
    # a bias that puts one pooled conv2 pre-activation exactly on the kink drives the conv2 gap to ~0
    pp = {k: v.clone() for k, v in params.items()}
    a = torch.nn.functional.max_pool2d(torch.relu(torch.nn.functional.conv2d(x.double(), pp["conv1.weight"].double(),
                                                                             pp["conv1.bias"].double())), 2, 2)
    pre = torch.nn.functional.max_pool2d(torch.nn.functional.conv2d(a, pp["conv2.weight"].double(),
                                                                    pp["conv2.bias"].double()), 2, 2)
    pp["conv2.bias"][0] -= pre[0, 0, 0, 0].float()
    assert mo.relu_kink_gap(pp, x)[0] < 1e-6


def test_c_oracle_mask_override_is_neutral_on_its_own_activations_and_flips_on_request():
    """c_oracle.loss_and_grads(mask_override=...) (used by the large-batch GPU parity test): forcing the oracle's own
    activations changes nothing; forcing one dead fc1 unit alive (a tiny positive value) opens exactly that gradient
    path and is reported."""
    import numpy as np
    from oracle import c_oracle
    params = {k: v.numpy() for k, v in mo.init_params(1).items()}
    x, t = mo.synthetic_batch(3, 17)
    act = c_oracle.forward(params, x.numpy())
    base = c_oracle.loss_and_grads(params, x.numpy(), t.numpy())
    same = c_oracle.loss_and_grads(params, x.numpy(), t.numpy(), mask_override=(act["p1"], act["p2"], act["h"]))
    assert same[3] == {"p1": (0, 0.0), "p2": (0, 0.0), "h": (0, 0.0)}
    assert all(np.array_equal(base[2][k], same[2][k]) for k in base[2])
    h = act["h"].copy()
    dead = np.argwhere(h == 0)[0]
    h[tuple(dead)] = 1e-6
    flipped = c_oracle.loss_and_grads(params, x.numpy(), t.numpy(), mask_override=(act["p1"], act["p2"], h))
    assert flipped[3]["h"][0] == 1 and abs(flipped[3]["h"][1] - 1e-6) < 1e-12
    assert not np.array_equal(base[2]["fc1.bias"], flipped[2]["fc1.bias"])
    assert np.array_equal(base[2]["fc2.bias"], flipped[2]["fc2.bias"])
