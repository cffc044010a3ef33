"""``DistributedDataParallel`` -- what the reference gets from
``nn.parallel.DistributedDataParallel`` / ``DistributedDataParallelCPU`` (mnist.py:135-138):
rank 0's parameters are broadcast at construction and every backward leaves the cross-rank
gradient in place.  Two exchange paths, both over ONE flat 1.7 MB gradient bucket:

* fused (default for 2/4/8 ranks on one NVLink-connected box): every rank owns a *peer window* -- one
  device allocation holding its parameter bucket, gradient-sum bucket (p.grad), two staging gradient
  buckets and a receive area -- that all other ranks map through CUDA IPC (``b200mnist_peer_window_*``; this
  works with one process per GPU under ``CUDA_VISIBLE_DEVICES=<rank>``, the single-GPU-per-pod view, as well as
  with all GPUs visible).  ``b200mnist_peer_train_step`` is then the whole step on one stream with the exchange
  inside its kernels: fc1's 93 % of the bucket is reduced two-shot (each rank sums its 1/world slice in rank
  order, applies SGD and stores the new parameters into every window) by extra CTAs in the launch of conv1's
  weight-gradient kernel; the small remainder is pushed to all peers and summed by the step's last kernel.  No
  NCCL call, no separate optimizer launch, no extra kernel.  fc1's momentum is kept by each slice's owner only
  (``flat_momentum`` holds this rank's slice of it).
* NCCL: ``all_reduce(SUM)`` of the bucket on the process group, then the fused SGD kernel with
  ``grad_scale = 1/world_size`` (also what the autograd-bridge path and gloo/CPU host tests use).

Which path a job takes is decided collectively at construction (every rank takes the same one) and
logged once on rank 0 with the reason.
"""
from __future__ import annotations

import ctypes
import logging

import torch
import torch.distributed as dist
import torch.nn as nn

from . import _cabi
from .net import Net, _stream

log = logging.getLogger("pytorch_operator_b200.ddp")
_IPC_BYTES = 64
_FLAG_BYTES = 4096
_PROBE_WORD = 192  # include/b200mnist.h: B200MNIST_PEER_PROBE_WORD


class _DeviceMemory:
    """Zero-copy torch view of raw device memory (``__cuda_array_interface__``)."""

    def __init__(self, ptr: int, count: int, typestr: str):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": typestr, "data": (ptr, False), "version": 3,
                                         "strides": None}


def _view(ptr: int, count: int, dtype, device) -> torch.Tensor:
    """A tensor over memory of THIS rank's own window (never a peer's mapping)."""
    typestr = {torch.float32: "<f4", torch.int32: "<i4", torch.int64: "<i8"}[dtype]
    t = torch.as_tensor(_DeviceMemory(ptr, count, typestr), device=device)
    if t.data_ptr() != ptr or t.dtype != dtype:
        raise RuntimeError("could not view the peer window as a tensor without a copy")
    return t


class DistributedDataParallel(nn.Module):
    def __init__(self, module: Net, process_group=None, fused=None):
        super().__init__()
        if not isinstance(module, Net):
            raise TypeError("this DistributedDataParallel wraps pytorch_operator_b200.Net only")
        if not dist.is_initialized():
            raise RuntimeError("init_process_group() first (mnist.py:116)")
        self.module = module
        self.process_group = process_group
        self.world_size = dist.get_world_size(process_group)
        self.rank = dist.get_rank(process_group)
        object.__setattr__(module, "_ddp", self)  # plain back-reference, not a registered sub-module (no cycle)
        dist.broadcast(module.flat_params, src=0, group=process_group)  # DDP ctor semantics
        self.fused = False
        self.fused_error = None
        self._windows = None
        self._opened = []
        self._step = 0              # steps taken on the fused path: step & 1 selects the staging bucket
        self._force_parity = None   # set while a StepGraph captures one of its two variants
        want = fused if fused is not None else True
        eligible = (want and module.device.type == "cuda" and self.world_size in (2, 4, 8)
                    and dist.get_backend(process_group) == "nccl")
        if eligible:
            self._setup_fused()
        elif want and fused:
            self.fused_error = "fused exchange needs CUDA, the nccl backend and 2, 4 or 8 ranks"
        if self.rank == 0:
            if self.fused:
                log.info("DDP exchange: fused peer-window kernels over NVLink (CUDA IPC), %d ranks", self.world_size)
            else:
                log.info("DDP exchange: %s all_reduce + SGD kernel%s", dist.get_backend(process_group),
                         "" if self.fused_error is None else " (fused path unavailable: %s)" % self.fused_error)
        if fused and not self.fused:
            raise RuntimeError("fused DDP exchange unavailable: %s" % self.fused_error)

    # ---- fused set-up: every decision that can differ between ranks is agreed on before the next step ----
    def _agree(self, ok: bool, err):
        """all ranks succeed or none does; returns the agreed outcome (and keeps the first local error text)."""
        if not ok and self.fused_error is None:
            self.fused_error = repr(err)
        flag = torch.tensor([1 if ok else 0], device=self.module.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.process_group)
        agreed = bool(flag.item())
        if not agreed and self.fused_error is None:
            self.fused_error = "another rank could not set up its peer window"
        return agreed

    def _setup_fused(self):
        net, ws, L = self.module, self.world_size, _cabi.lib()
        dev = net.device
        own, handle, err = ctypes.c_void_p(), (ctypes.c_char * _IPC_BYTES)(), None
        try:  # 1. local: allocate the window and export it
            with torch.cuda.device(dev):
                _cabi.check(L.b200mnist_peer_window_create(ctypes.byref(own), handle), "peer_window_create")
        except Exception as e:
            err = e
        if not self._agree(err is None, err):
            if own.value:
                L.b200mnist_peer_window_destroy(own)
            return
        # 2. exchange the handles over the process group (the env:// rendezvous the operator set up) and map the peers
        handles = [None] * ws
        dist.all_gather_object(handles, bytes(handle.raw), group=self.process_group)
        ptrs, opened = [0] * ws, []
        try:
            with torch.cuda.device(dev):
                for r in range(ws):
                    if r == self.rank:
                        ptrs[r] = own.value
                        continue
                    p = ctypes.c_void_p()
                    _cabi.check(L.b200mnist_peer_window_open(handles[r], ctypes.byref(p)), f"peer_window_open(rank {r})")
                    ptrs[r] = p.value
                    opened.append(p.value)
            # 3. verify each mapping: every rank stamps its own window, then reads all stamps through the mappings
            _view(own.value + 4 * _PROBE_WORD, 1, torch.int32, dev).fill_(1000 + self.rank)
            torch.cuda.synchronize(dev)
        except Exception as e:
            err = e
        dist.barrier(group=self.process_group)
        if err is None:
            try:
                got = torch.zeros(ws, dtype=torch.int32, device=dev)
                _cabi.check(L.b200mnist_peer_window_probe((ctypes.c_void_p * ws)(*ptrs), ws, got.data_ptr(), _stream()),
                            "peer_window_probe")
                got = got.tolist()
                if got != [1000 + r for r in range(ws)]:
                    raise RuntimeError(f"peer windows read back {got}")
            except Exception as e:
                err = e
        if not self._agree(err is None, err):
            with torch.cuda.device(dev):
                for p in opened:
                    L.b200mnist_peer_window_close(p)
                dist.barrier(group=self.process_group)
                L.b200mnist_peer_window_destroy(own)
            return
        # 4. move the model into the window: parameters and p.grad become views of window memory
        n = net.flat_params.numel()
        bucket = 4 * n
        assert int(L.b200mnist_peer_window_bytes()) >= _FLAG_BYTES + 4 * bucket
        params = _view(own.value + _FLAG_BYTES, n, torch.float32, dev)
        gsum = _view(own.value + _FLAG_BYTES + bucket, n, torch.float32, dev)
        params.copy_(net.flat_params)
        gsum.copy_(net.flat_grads)
        net._bind(params, gsum, net.flat_momentum)
        self._staging = [_view(own.value + _FLAG_BYTES + (2 + k) * bucket, n, torch.float32, dev) for k in (0, 1)]
        self._stamps = _view(own.value + 2048, 8, torch.int64, dev)   # peer_exchange.cuh: kStampByteOffset
        self._windows = (ctypes.c_void_p * ws)(*ptrs)
        self._own_window, self._opened = own, opened
        torch.cuda.synchronize(dev)
        dist.barrier(group=self.process_group)
        self.fused = True

    def close(self):
        """Unmap the peers' windows and free this rank's (collective; call before destroy_process_group()).  The
        model's buckets move back to ordinary device memory first, so the model stays usable (NCCL path)."""
        if not self.fused:
            return
        net, L = self.module, _cabi.lib()
        net.release_graphs()
        torch.cuda.synchronize(net.device)
        net._bind(net.flat_params.clone(), net.flat_grads.clone(), net.flat_momentum)
        self.fused, self._staging, self._windows = False, None, None
        dist.barrier(group=self.process_group)
        with torch.cuda.device(net.device):
            for p in self._opened:
                L.b200mnist_peer_window_close(p)
            dist.barrier(group=self.process_group)
            L.b200mnist_peer_window_destroy(self._own_window)
        self._opened = []

    # ---- exchange paths ----
    def reduce_gradients(self):
        """NCCL path: sum the flat gradient bucket over all ranks, in place, on the current stream."""
        dist.all_reduce(self.module.flat_grads, op=dist.ReduceOp.SUM, group=self.process_group)

    def fused_train_step(self, x, target, optimizer, loss_out=None, ring=None):
        """Fused path: the whole step, exchange included, as one C call on the current stream
        (b200mnist_peer_train_step; with ``ring`` the batch comes from this rank's resident InputRing)."""
        net = self.module
        capturing = self._force_parity is not None
        parity = self._force_parity if capturing else self._step & 1
        B = x.shape[0]
        ws = net.workspace(B)
        lo = net._loss if loss_out is None else loss_out
        L = _cabi.lib()
        head = (self._windows, self.rank, self.world_size, parity, net.flat_momentum.data_ptr())
        tail = (x.data_ptr(), target.data_ptr(), B, ws.data_ptr(), ws.numel(), optimizer.lr, optimizer.momentum,
                1 if optimizer.first_step else 0, lo.data_ptr(), _stream())
        if ring is None:
            _cabi.check(L.b200mnist_peer_train_step(*head, *tail), "peer_train_step")
        else:
            _cabi.check(L.b200mnist_peer_train_step_ring(*head, ring.desc_ptr, *tail), "peer_train_step_ring")
        optimizer.first_step = False
        if not capturing:
            self._step += 1
        return lo[0]

    def exchange_telemetry(self):
        """Where the latest fused step's exchange sat in time (microseconds after the start of the exchange CTAs in
        conv1's weight-gradient kernel; GPU globaltimer stamps the kernels leave in the flag page):
        {fc1_barrier, fc1_slice_done, tail_start, tail_pushed, tail_summed, step_period}.  Synchronises the device."""
        if not self.fused:
            return None
        torch.cuda.synchronize(self.module.device)
        s = self._stamps.tolist()
        rel = lambda k: round((s[k] - s[0]) * 1e-3, 2)
        return {"fc1_barrier": rel(1), "fc1_slice_done": rel(2), "tail_start": rel(3), "tail_pushed": rel(4),
                "tail_summed": rel(5), "step_period": round((s[5] - s[6]) * 1e-3, 2)}

    @property
    def staging(self):
        """The staging bucket the NEXT fused step writes this rank's gradient into (None on the NCCL path)."""
        return self._staging[self._step & 1] if self.fused else None

    def forward(self, x):
        return self.module(x)

    def train_step(self, x, target, optimizer, loss_out=None, ring=None):
        return self.module.train_step(x, target, optimizer, loss_out=loss_out, ring=ring)

    def evaluate(self, *a, **k):
        return self.module.evaluate(*a, **k)

    def release_graphs(self):
        self.module.release_graphs()

    @property
    def device(self):
        return self.module.device
