"""``DistributedDataParallel`` -- what the reference gets from
``nn.parallel.DistributedDataParallel`` / ``DistributedDataParallelCPU`` (mnist.py:135-138):
rank 0's parameters are broadcast at construction and every backward leaves the cross-rank
gradient in place.  Two exchange paths, both over ONE flat 1.7 MB gradient bucket:

* fused (default for 2/4/8 ranks on one NVLink-connected box): every rank owns a *peer window* -- one
  device allocation holding its parameter bucket, gradient-sum bucket (p.grad) and two staging gradient
  buckets -- that all other ranks map through CUDA IPC (``b200mnist_peer_window_*``; this works with one
  process per GPU under ``CUDA_VISIBLE_DEVICES=<rank>``, the single-GPU-per-pod view, as well as with all
  GPUs visible).  ``b200mnist_peer_exchange`` then replaces all_reduce + optimizer.step(): fc1's 93 % of the
  bucket is reduced two-shot (each rank sums its 1/world slice in rank order, applies SGD and stores the new
  parameters into every window) on a side stream WHILE the conv backward runs; the small remainder is
  reduced one-shot at the end of the step.  No NCCL call, no separate optimizer launch.  fc1's momentum is
  kept by each slice's owner only (``flat_momentum`` holds this rank's slice of it).
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
    typestr = {torch.float32: "<f4", torch.int32: "<i4"}[dtype]
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
        self.timeline = None        # set to a dict to have the next eager fused step record its phase boundaries
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
        assert int(L.b200mnist_peer_window_bytes()) == _FLAG_BYTES + 4 * bucket
        params = _view(own.value + _FLAG_BYTES, n, torch.float32, dev)
        gsum = _view(own.value + _FLAG_BYTES + bucket, n, torch.float32, dev)
        params.copy_(net.flat_params)
        gsum.copy_(net.flat_grads)
        net._bind(params, gsum, net.flat_momentum)
        self._staging = [_view(own.value + _FLAG_BYTES + (2 + k) * bucket, n, torch.float32, dev) for k in (0, 1)]
        self._windows = (ctypes.c_void_p * ws)(*ptrs)
        self._own_window, self._opened = own, opened
        self._side = torch.cuda.Stream(device=dev)
        self._ev_fc, self._ev_side = torch.cuda.Event(), torch.cuda.Event()
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

    def _exchange(self, optimizer, parity, part):
        m = self.module
        _cabi.check(_cabi.lib().b200mnist_peer_exchange(
            self._windows, self.rank, self.world_size, parity, m.flat_momentum.data_ptr(), optimizer.lr,
            optimizer.momentum, 1.0 / self.world_size, 1 if optimizer.first_step else 0, part, _stream()),
            "peer_exchange")

    def fused_train_step(self, x, target, optimizer, loss_out=None):
        """Fused path.  The gradient is produced into this step's staging bucket in two phases; fc1's part
        (93 % of the bucket, final after the fc backward kernel) is exchanged + applied by the two-shot peer kernel
        on a side stream WHILE the conv backward runs, the remainder right after it."""
        net = self.module
        cur = torch.cuda.current_stream()
        capturing = self._force_parity is not None
        parity = self._force_parity if capturing else self._step & 1
        staging = self._staging[parity]
        tl = None if capturing else self.timeline

        def mark(key, stream):
            if tl is not None:
                tl[key] = torch.cuda.Event(enable_timing=True)
                tl[key].record(stream)

        mark("start", cur)
        net.forward_backward_phase(x, target, 1, staging, loss_out)
        mark("phase1_end", cur)
        self._ev_fc.record(cur)
        self._side.wait_event(self._ev_fc)
        with torch.cuda.stream(self._side):
            self._exchange(optimizer, parity, 1)
            mark("fc1_exchange_end", self._side)
            self._ev_side.record(self._side)
        loss = net.forward_backward_phase(x, target, 2, staging, loss_out)
        mark("phase2_end", cur)
        self._exchange(optimizer, parity, 2)
        mark("rest_exchange_end", cur)
        cur.wait_event(self._ev_side)
        optimizer.first_step = False
        if not capturing:
            self._step += 1
        return loss

    @property
    def staging(self):
        """The staging bucket the NEXT fused step writes this rank's gradient into (None on the NCCL path)."""
        return self._staging[self._step & 1] if self.fused else None

    def forward(self, x):
        return self.module(x)

    def train_step(self, x, target, optimizer, loss_out=None):
        return self.module.train_step(x, target, optimizer, loss_out=loss_out)

    def evaluate(self, *a, **k):
        return self.module.evaluate(*a, **k)

    def release_graphs(self):
        self.module.release_graphs()

    @property
    def device(self):
        return self.module.device
