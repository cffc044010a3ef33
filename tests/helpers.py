"""Shared helpers for the parity tests."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NAMES = ["conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias",
         "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias"]
# fp32 parity tolerance north_star states: 1e-5 relative on per-step loss and gradients.
# Gradients contain exact zeros and near-zero entries (dead ReLU columns), where a pure
# relative bound is meaningless, so each tensor gets atol = RTOL * max|reference tensor|.
RTOL = 1e-5


def load_golden(name):
    with np.load(os.path.join(GOLDEN, name + ".npz")) as z:
        return {k: z[k] for k in z.files}


def golden_tensor_check(gold, prefix, name, value, rtol=RTOL, atol_scale=1.0):
    """Compare a full tensor `value` with what the golden file stores under prefix/name
    (whole tensor, or stride sample + sum/l2 for fc1.weight)."""
    value = np.asarray(value, dtype=np.float32)
    key = f"{prefix}/{name}"
    if key in gold:
        ref = gold[key]
        assert_close(value.reshape(ref.shape), ref, rtol, atol_scale, what=key)
    else:
        stride = int(gold["meta/fc1_stride"][0])
        ref = gold[key + "/sample"]
        assert_close(value.reshape(-1)[::stride], ref, rtol, atol_scale, what=key + "/sample")
        s, l2 = gold[key + "/sum_l2"]
        v64 = value.reshape(-1).astype(np.float64)
        assert abs(np.sqrt((v64 ** 2).sum()) - l2) <= 1e-5 * l2 + 1e-12, key + " l2"


def assert_close(a, b, rtol=RTOL, atol_scale=1.0, what=""):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    scale = float(np.max(np.abs(b))) if b.size else 0.0
    atol = rtol * scale * atol_scale
    err = np.abs(a - b)
    bound = atol + rtol * np.abs(b)
    bad = err > bound
    assert not bad.any(), (f"{what}: {int(bad.sum())}/{a.size} out of tolerance; max err "
                           f"{float(err.max()):.3e} (max|ref| {scale:.3e}, rtol {rtol})")


def tie_free_batch(params, B, seed, min_gap=5e-7):
    """synthetic_batch(B, seed'), with seed' the first seed >= `seed` whose batch has no
    max-pool near-tie under `params` (oracle.mnist_oracle.pool_tie_gap)."""
    from oracle import mnist_oracle as mo
    for s in range(seed, seed + 2000):
        x, t = mo.synthetic_batch(B, s)
        if mo.pool_tie_gap(params, x) >= min_gap:
            return x, t
    raise AssertionError("no tie-free batch found")


def kink_free_batch(params, B, seed, min_conv2=5e-6, min_fc1=0.0):
    """synthetic_batch(B, seed'), with seed' the first seed >= `seed` whose pooled conv2 (and, if asked, fc1)
    pre-activations all stay that far away from zero (oracle.mnist_oracle.relu_kink_gap): the batches on which the
    3xTF32 kernels and fp32 ones must agree on every ReLU mask."""
    from oracle import mnist_oracle as mo
    for s in range(seed, seed + 400):
        x, t = mo.synthetic_batch(B, s)
        g2, g1 = mo.relu_kink_gap(params, x)
        if g2 >= min_conv2 and g1 >= min_fc1:
            return x, t
    raise AssertionError("no kink-free batch found")


def golden_batch(gold, B, step, rank):
    """The batch gen_golden.py used for (step, rank): the seed is stored in the fixture."""
    from oracle import mnist_oracle as mo
    return mo.synthetic_batch(B, int(gold[f"step{step}/seed_rank{rank}"][0]))
