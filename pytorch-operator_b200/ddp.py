"""``DistributedDataParallel`` -- what the reference gets from
``nn.parallel.DistributedDataParallel`` / ``DistributedDataParallelCPU`` (mnist.py:135-138):
rank 0's parameters are broadcast at construction and every backward leaves the cross-rank
gradient in place.  Two exchange paths, both over ONE flat 1.7 MB gradient bucket:

* fused (default on one NVLink-connected box): the step's gradient is written into a
  symmetric-memory bucket every peer has mapped, and ``b200mnist_allreduce_sgd`` -- one kernel --
  reads all ranks' buckets over NVLink, sums them in rank order, and applies the SGD update with
  1/world_size folded in (no NCCL call, no separate optimizer launch);
* NCCL: ``all_reduce(SUM)`` of the bucket on the process group, then the fused SGD kernel with
  ``grad_scale = 1/world_size`` (also what the autograd-bridge path and gloo/CPU host tests use).
"""
from __future__ import annotations

import ctypes

import torch
import torch.distributed as dist
import torch.nn as nn

from . import _cabi
from .net import Net, _stream

_PAD_OFFSET_BYTES = 8192  # our flags live at the end of the 9 KiB signal pad, away from torch's own channels


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
        want = fused if fused is not None else True
        if want and module.device.type == "cuda" and 2 <= self.world_size <= 8 \
                and dist.get_backend(process_group) == "nccl":
            try:
                self._setup_fused()
                self.fused = True
            except Exception as e:  # no peer access / symmetric memory unavailable: NCCL path
                self.fused_error = repr(e)
                if fused:
                    raise
        # every rank must take the same path
        flag = torch.tensor([1 if self.fused else 0], device=module.device if module.device.type == "cuda" else "cpu")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=process_group)
        self.fused = bool(flag.item())

    def _setup_fused(self):
        import torch.distributed._symmetric_memory as symm
        net, ws = self.module, self.world_size
        group = self.process_group if self.process_group is not None else dist.group.WORLD
        self._staging = symm.empty(net.flat_grads.numel(), dtype=torch.float32, device=net.device)
        self._staging.zero_()
        self._symm = symm.rendezvous(self._staging, group)
        if int(getattr(self._symm, "signal_pad_size", 0)) < _PAD_OFFSET_BYTES + 256:
            raise RuntimeError("signal pad too small")
        self._peer_grads = (ctypes.c_void_p * ws)(*[int(p) for p in self._symm.buffer_ptrs])
        self._peer_signals = (ctypes.c_void_p * ws)(*[int(p) + _PAD_OFFSET_BYTES for p in self._symm.signal_pad_ptrs])
        self._state = torch.zeros(4, dtype=torch.int32, device=net.device)   # two exchange instances x (epoch, ticket)
        self._peer_signals_b = (ctypes.c_void_p * ws)(*[int(p) + _PAD_OFFSET_BYTES + 256 for p in self._symm.signal_pad_ptrs])
        L = _cabi.lib()
        self._fc1_range = (int(L.b200mnist_tensor_offset(4)), int(L.b200mnist_tensor_offset(6)))
        self._side = torch.cuda.Stream(device=net.device)
        self._ev_fc, self._ev_side = torch.cuda.Event(), torch.cuda.Event()
        torch.cuda.synchronize(net.device)
        dist.barrier(group=self.process_group)

    # ---- exchange paths ----
    def reduce_gradients(self):
        """NCCL path: sum the flat gradient bucket over all ranks, in place, on the current stream."""
        dist.all_reduce(self.module.flat_grads, op=dist.ReduceOp.SUM, group=self.process_group)

    def _allreduce_sgd(self, optimizer, signals, state_ptr, only_fc1):
        m = self.module
        _cabi.check(_cabi.lib().b200mnist_allreduce_sgd(
            m.flat_params.data_ptr(), m.flat_momentum.data_ptr(), m.flat_grads.data_ptr(), self._peer_grads,
            signals, self.rank, self.world_size, state_ptr, optimizer.lr, optimizer.momentum,
            1.0 / self.world_size, 1 if optimizer.first_step else 0, self._fc1_range[0], self._fc1_range[1],
            1 if only_fc1 else 0, _stream()), "allreduce_sgd")

    def fused_train_step(self, x, target, optimizer, loss_out=None):
        """Fused path.  The gradient is produced into the symmetric staging bucket in two phases; fc1's part
        (93 % of the bucket, final after the fc backward kernel) is exchanged + applied by the peer-memory kernel
        on a side stream WHILE the conv backward runs, the remainder right after it.  Peers' staged gradients ->
        rank-ordered sum in flat_grads -> SGD update, 1/world_size folded in."""
        net = self.module
        cur = torch.cuda.current_stream()
        net.forward_backward_phase(x, target, 1, self._staging, loss_out)
        self._ev_fc.record(cur)
        self._side.wait_event(self._ev_fc)
        with torch.cuda.stream(self._side):
            self._allreduce_sgd(optimizer, self._peer_signals, self._state.data_ptr(), only_fc1=True)
            self._ev_side.record(self._side)
        loss = net.forward_backward_phase(x, target, 2, self._staging, loss_out)
        self._allreduce_sgd(optimizer, self._peer_signals_b, self._state.data_ptr() + 8, only_fc1=False)
        cur.wait_event(self._ev_side)
        optimizer.first_step = False
        return loss

    @property
    def staging(self):
        return self._staging if self.fused else None

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
