"""The allocation-free head entry point, b200mnist_fc2_logsoftmax_nll_static.

It exists for hosts that cannot give the library a scratch buffer because they are not allowed
to allocate device memory at the point of the call (a callback inside someone else's runtime,
where cudaMalloc could stall the device; a process whose GPU memory belongs to a foreign pool).
Its scratch is a fixed array inside the library, shared by every call and never cleared
between them, so this test makes two calls in a row -- with different batches -- and checks each
against the ordinary entry point (caller-provided scratch) run on the same inputs.

This file sorts last on purpose: a run that stops at the first failure has by then run every
other test.  It also passes on its own in a fresh process (pytest tests/test_zz_no_alloc.py -m gpu).
"""
import numpy as np
import pytest
import torch

from tests.helpers import RTOL, assert_close

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _inputs(B, seed):
    g = torch.Generator().manual_seed(seed)
    h = torch.relu(torch.randn(B, 500, generator=g)).to(DEV)
    w = (torch.randn(10, 500, generator=g) * 0.05).to(DEV)
    b = (torch.randn(10, generator=g) * 0.05).to(DEV)
    t = torch.randint(0, 10, (B,), generator=g).to(DEV)
    return h, w, b, t


def _call(static, h, w, b, t):
    from pytorch_operator_b200 import _cabi
    L = _cabi.lib()
    B = h.shape[0]
    out = {"logp": torch.full((B, 10), 7.0, device=DEV), "loss": torch.full((1,), 7.0, device=DEV),
           "dz2": torch.full((B, 10), 7.0, device=DEV), "dz1": torch.full((B, 500), 7.0, device=DEV)}
    stream = torch.cuda.current_stream().cuda_stream
    if static:
        rc = L.b200mnist_fc2_logsoftmax_nll_static(h.data_ptr(), w.data_ptr(), b.data_ptr(), t.data_ptr(), B,
                                                   out["logp"].data_ptr(), out["loss"].data_ptr(),
                                                   out["dz2"].data_ptr(), out["dz1"].data_ptr(), stream)
    else:
        scratch = torch.zeros(B + 64, device=DEV)
        rc = L.b200mnist_fc2_logsoftmax_nll(h.data_ptr(), w.data_ptr(), b.data_ptr(), t.data_ptr(), B,
                                            out["logp"].data_ptr(), out["loss"].data_ptr(), out["dz2"].data_ptr(),
                                            out["dz1"].data_ptr(), scratch.data_ptr(), stream)
    _cabi.check(rc, "head entry point")
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in out.items()}


def test_static_scratch_entry_point_matches_the_regular_one_twice_in_a_row():
    # two different batches back to back: the second call finds the first call's leftovers
    # (per-row losses of a LARGER batch, a used ticket) in the shared array
    for B, seed in ((200, 21), (37, 22)):
        h, w, b, t = _inputs(B, seed)
        got = _call(True, h, w, b, t)
        want = _call(False, h, w, b, t)
        for k in want:
            assert_close(got[k], want[k], rtol=RTOL, what=f"{k} (B={B})")
        lp = torch.log_softmax(h.cpu() @ w.cpu().T + b.cpu(), dim=1)  # and against plain torch on the CPU
        assert_close(got["logp"], lp.numpy(), what="logp vs torch")
        assert abs(float(got["loss"][0]) - float(torch.nn.functional.nll_loss(lp, t.cpu()))) <= 2 * RTOL * 2.4


def test_static_scratch_entry_point_rejects_oversized_batches():
    from pytorch_operator_b200 import _cabi
    L = _cabi.lib()
    h, w, b, t = _inputs(8, 5)
    out = torch.zeros(1, device=DEV)
    rc = L.b200mnist_fc2_logsoftmax_nll_static(h.data_ptr(), w.data_ptr(), b.data_ptr(), t.data_ptr(), 4097, 0,
                                               out.data_ptr(), 0, 0, torch.cuda.current_stream().cuda_stream)
    assert rc == -10006
    assert "B200MNIST_STATIC_MAX_BATCH" in _cabi.last_error() and "4096" in _cabi.last_error()
    # the largest allowed batch works
    h, w, b, t = _inputs(4096, 6)
    got, want = _call(True, h, w, b, t), _call(False, h, w, b, t)
    assert np.array_equal(got["logp"], want["logp"])
    assert_close(got["loss"], want["loss"], what="loss at the size limit")
