"""ctypes view of oracle/liblenet_ref.so (the plain-C restatement, lenet_ref.c).

TEST INFRASTRUCTURE ONLY -- see the header of lenet_ref.c.  numpy in, numpy out.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liblenet_ref.so")
NAMES = ["conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias",
         "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias"]


def build(force=False):
    src = os.path.join(_HERE, "lenet_ref.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "liblenet_ref.so"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.lenet_nll_mean.restype = ctypes.c_double
    return _lib


def _f(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(ctypes.c_void_p)


def forward(params, x):
    """params: dict name->float32 array; x: [B,1,28,28].  Returns dict of activations."""
    B = int(x.shape[0])
    keep = [_f(params[n]) for n in NAMES]
    xa, xp = _f(x)
    act = {"p1": np.zeros((B, 20, 12, 12), np.float32), "i1": np.zeros((B, 20, 12, 12), np.uint8),
           "p2": np.zeros((B, 50, 4, 4), np.float32), "i2": np.zeros((B, 50, 4, 4), np.uint8),
           "h": np.zeros((B, 500), np.float32), "logp": np.zeros((B, 10), np.float32)}
    vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    lib().lenet_forward(*[k[1] for k in keep], xp, ctypes.c_int(B), vp(act["p1"]), vp(act["i1"]),
                        vp(act["p2"]), vp(act["i2"]), vp(act["h"]), vp(act["logp"]))
    act["x"] = xa
    return act


def loss_and_grads(params, x, target, idx_override=None, mask_override=None):
    """Returns (loss, logp, grads dict) for one batch; the C analogue of
    mnist_oracle.loss_and_grads.  ``idx_override=(i1, i2)`` teacher-forces the max-pool
    arg-max maps used by the backward (see mnist_oracle.pool_tie_gap).
    ``mask_override=(p1, p2, h)`` (the activations another implementation saved) teacher-forces the
    three ReLU masks of the backward as well: wherever that implementation and this one disagree on
    ``activation > 0`` the other one's value is taken; a fourth return value then reports, per layer,
    how many masks were forced and the largest activation magnitude among them (a legitimate
    disagreement is one within rounding of the kink, i.e. that magnitude is ~1e-6)."""
    B = int(x.shape[0])
    act = forward(params, x)
    if idx_override is not None:
        act["i1"] = np.ascontiguousarray(idx_override[0], dtype=np.uint8).reshape(act["i1"].shape)
        act["i2"] = np.ascontiguousarray(idx_override[1], dtype=np.uint8).reshape(act["i2"].shape)
    forced = None
    if mask_override is not None:
        forced = {}
        for key, other in zip(("p1", "p2", "h"), mask_override):
            mine = act[key]
            other = np.asarray(other, dtype=np.float32).reshape(mine.shape)
            diff = (mine > 0) != (other > 0)
            mag = max(float(np.abs(mine[diff]).max(initial=0.0)), float(np.abs(other[diff]).max(initial=0.0)))
            forced[key] = (int(diff.sum()), mag)
            mine[diff] = other[diff]
    t = np.ascontiguousarray(target, dtype=np.int64)
    vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    loss = lib().lenet_nll_mean(vp(act["logp"]), vp(t), ctypes.c_int(B))
    grads = {n: np.zeros(np.shape(params[n]), np.float32) for n in NAMES}
    keep = {n: _f(params[n]) for n in ("conv2.weight", "fc1.weight", "fc2.weight")}
    lib().lenet_backward(keep["conv2.weight"][1], keep["fc1.weight"][1], keep["fc2.weight"][1],
                         vp(act["x"]), vp(t), ctypes.c_int(B), vp(act["p1"]), vp(act["i1"]), vp(act["p2"]),
                         vp(act["i2"]), vp(act["h"]), vp(act["logp"]), *[vp(grads[n]) for n in NAMES])
    if forced is not None:
        return float(loss), act["logp"], grads, forced
    return float(loss), act["logp"], grads


def sgd_momentum_(p, g, buf, lr, mu, grad_scale, first_step):
    """In-place on float32 numpy arrays p and buf."""
    assert p.dtype == np.float32 and buf.dtype == np.float32 and p.flags.c_contiguous and buf.flags.c_contiguous
    ga, gp = _f(g)
    vp = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    lib().lenet_sgd_momentum(vp(p), gp, vp(buf), ctypes.c_long(p.size), ctypes.c_float(lr), ctypes.c_float(mu),
                             ctypes.c_float(grad_scale), ctypes.c_int(1 if first_step else 0))


def square(a):
    aa, ap = _f(a)
    out = np.empty_like(aa)
    lib().smoke_square(ap, out.ctypes.data_as(ctypes.c_void_p), ctypes.c_long(aa.size))
    return out
