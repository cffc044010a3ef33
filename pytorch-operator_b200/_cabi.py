"""ctypes binding of libb200mnist.so (C ABI declared in include/b200mnist.h).

There is no CPU fallback: if the library is missing and cannot be built, or a call
fails, a RuntimeError is raised.  ``build()`` compiles the library in-tree with nvcc
for sm_100a; on the GPU box the prebuilt .so travels with the repo snapshot.
"""
from __future__ import annotations

import ctypes
import os
import shutil
import subprocess

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
_SRC_DIR = os.path.join(_PKG, "csrc")
SO_PATH = os.path.join(_PKG, "libb200mnist.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC"]


def _sources():
    out = [os.path.join(_ROOT, "include", "b200mnist.h")]
    for f in sorted(os.listdir(_SRC_DIR)):
        if f.endswith((".cu", ".cuh")):
            out.append(os.path.join(_SRC_DIR, f))
    return out


def _stale() -> bool:
    if not os.path.exists(SO_PATH):
        return True
    t = os.path.getmtime(SO_PATH)
    return any(os.path.getmtime(s) > t for s in _sources())


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/b200mnist.cu -> libb200mnist.so (sm_100a) if missing or out of date."""
    if not force and not _stale():
        return SO_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        if os.path.exists(SO_PATH):
            return SO_PATH  # sources look newer (e.g. fresh checkout mtimes) but no compiler: use what travelled
        raise RuntimeError("libb200mnist.so is missing and nvcc is not available to build it")
    tmp = f"{SO_PATH}.{os.getpid()}.tmp"  # per process: several ranks may find the library stale at the same time
    cmd = [nvcc, *NVCC_FLAGS, "-o", tmp, os.path.join(_SRC_DIR, "b200mnist.cu")]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        if os.path.exists(tmp):
            os.unlink(tmp)
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    os.replace(tmp, SO_PATH)  # atomic: a concurrent loader sees the old or the new file, never a partial one
    if verbose:
        print(res.stderr)
    return SO_PATH


_c_f = ctypes.c_void_p  # device pointers travel as raw addresses
_i32, _i64, _f32 = ctypes.c_int, ctypes.c_int64, ctypes.c_float

# name -> (restype, argtypes); must list every symbol include/b200mnist.h declares
SIGNATURES = {
    "b200mnist_flat_elems": (_i64, []),
    "b200mnist_tensor_offset": (_i64, [_i32]),
    "b200mnist_tensor_elems": (_i64, [_i32]),
    "b200mnist_last_error": (ctypes.c_char_p, []),
    "b200mnist_version": (_i32, []),
    "b200mnist_workspace_bytes": (_i64, [_i32]),
    "b200mnist_workspace_offset": (_i64, [_i32, _i32]),
    "b200mnist_forward_backward": (_i32, [_c_f, _c_f, _c_f, _c_f, _i32, _c_f, _i64, _c_f, _c_f, _c_f]),
    "b200mnist_train_step": (_i32, [_c_f, _c_f, _c_f, _c_f, _c_f, _i32, _c_f, _i64, _f32, _f32, _i32, _c_f, _c_f]),
    "b200mnist_train_step_ring": (_i32, [_c_f, _c_f, _c_f, _c_f, _c_f, _c_f, _i32, _c_f, _i64, _f32, _f32, _i32, _c_f, _c_f]),
    "b200mnist_forward_eval": (_i32, [_c_f, _c_f, _c_f, _i32, _c_f, _i64, _c_f, _c_f, _c_f]),
    "b200mnist_sgd_momentum": (_i32, [_c_f, _c_f, _c_f, _i64, _f32, _f32, _f32, _i32, _c_f]),
    "b200mnist_peer_train_step_ring": (_i32, [ctypes.POINTER(ctypes.c_void_p), _i32, _i32, _i32, _c_f, _c_f, _c_f, _c_f,
                                              _i32, _c_f, _i64, _f32, _f32, _i32, _c_f, _c_f]),
    "b200mnist_peer_window_bytes": (_i64, []),
    "b200mnist_peer_window_create": (_i32, [ctypes.POINTER(ctypes.c_void_p), _c_f]),
    "b200mnist_peer_window_open": (_i32, [_c_f, ctypes.POINTER(ctypes.c_void_p)]),
    "b200mnist_peer_window_close": (_i32, [_c_f]),
    "b200mnist_peer_window_destroy": (_i32, [_c_f]),
    "b200mnist_peer_window_probe": (_i32, [ctypes.POINTER(ctypes.c_void_p), _i32, _c_f, _c_f]),
    "b200mnist_peer_train_step": (_i32, [ctypes.POINTER(ctypes.c_void_p), _i32, _i32, _i32, _c_f, _c_f, _c_f, _i32, _c_f,
                                         _i64, _f32, _f32, _i32, _c_f, _c_f]),
    "b200mnist_conv1_relu_pool_fwd": (_i32, [_c_f, _c_f, _c_f, _c_f, _c_f, _i32, _c_f]),
    "b200mnist_conv2_relu_pool_fwd": (_i32, [_c_f, _c_f, _c_f, _c_f, _c_f, _i32, _c_f]),
    "b200mnist_conv2_fwd_tc_scratch_bytes": (_i64, []),
    "b200mnist_conv2_relu_pool_fwd_tc": (_i32, [_c_f, _c_f, _c_f, _c_f, _c_f, _i32, _c_f, _c_f]),
    "b200mnist_fc1_scratch_bytes": (_i64, [_i32]),
    "b200mnist_fc1_relu_fwd": (_i32, [_c_f, _c_f, _c_f, _c_f, _i32, _c_f, _c_f]),
    "b200mnist_fc1_relu_fwd_tc": (_i32, [_c_f, _c_f, _c_f, _c_f, _i32, _c_f, _c_f]),
    "b200mnist_fc2_logsoftmax_nll": (_i32, [_c_f, _c_f, _c_f, _c_f, _i32, _c_f, _c_f, _c_f, _c_f, _c_f, _c_f]),
    "b200mnist_fc2_logsoftmax_nll_static": (_i32, [_c_f, _c_f, _c_f, _c_f, _i32, _c_f, _c_f, _c_f, _c_f, _c_f]),
    "b200mnist_head_bwd": (_i32, [_c_f, _c_f, _c_f, _c_f, _i32, _c_f, _c_f, _c_f]),
    "b200mnist_fc_bwd_scratch_bytes": (_i64, [_i32]),
    "b200mnist_fc_bwd": (_i32, [_c_f, _c_f, _c_f, _c_f, _c_f, _i32, _c_f, _c_f, _c_f, _c_f, _c_f, _c_f, _c_f]),
    "b200mnist_conv2_bwd_scratch_bytes": (_i64, [_i32]),
    "b200mnist_conv2_bwd": (_i32, [_c_f, _c_f, _c_f, _c_f, _c_f, _i32, _c_f, _c_f, _c_f, _c_f, _c_f]),
    "b200mnist_conv1_bwd_scratch_bytes": (_i64, [_i32]),
    "b200mnist_conv1_bwd": (_i32, [_c_f, _c_f, _c_f, _c_f, _i32, _c_f, _c_f, _c_f, _c_f]),
    "b200mnist_gather_normalize_u8": (_i32, [_c_f, _c_f, _c_f, _i32, _i32, _f32, _f32, _c_f, _c_f, _c_f]),
    "b200mnist_square": (_i32, [_c_f, _c_f, _i64, _c_f]),
    "b200mnist_stage_inputs": (_i32, [_c_f, _c_f, _i64, _c_f, _c_f, _c_f, _c_f]),
    "b200mnist_event_record": (_i32, [_c_f, _c_f]),
    "b200mnist_stage_record": (_i32, [_c_f, _c_f, _i64, _c_f, _c_f, ctypes.c_uint32]),
    "b200mnist_launch_count": (ctypes.c_longlong, []),
    "b200mnist_delay": (_i32, [_c_f, _i32]),
    "b200mnist_fma_peak": (_i64, [_c_f, _i32, _i32, _c_f]),
    "b200mnist_profile_begin": (_i32, [_c_f, _i32]),
    "b200mnist_profile_end": (_i32, [ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_char_p), _i32, _c_f]),
}



class Ring(ctypes.Structure):
    """b200mnist_ring_t (include/b200mnist.h)."""
    _fields_ = [("x", ctypes.c_void_p), ("target", ctypes.c_void_p), ("x_stride", ctypes.c_int64),
                ("t_stride", ctypes.c_int64), ("slots", ctypes.c_uint32), ("state", ctypes.c_void_p),
                ("landed", ctypes.c_void_p), ("report_host", ctypes.c_void_p)]


_lib = None


def lib() -> ctypes.CDLL:
    """Load (building first if needed) and type the library.  Raises if unavailable."""
    global _lib
    if _lib is None:
        path = build()
        handle = ctypes.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError here = header/library mismatch
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def last_error() -> str:
    return (lib().b200mnist_last_error() or b"").decode()


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed (rc={rc}): {last_error()}")


def ptr(t) -> int:
    """Raw address of a torch tensor's first element (None -> NULL)."""
    return 0 if t is None else t.data_ptr()
