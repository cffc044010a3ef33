"""``DistributedDataParallel`` -- what the reference gets from
``nn.parallel.DistributedDataParallel`` / ``DistributedDataParallelCPU`` (mnist.py:135-138):
rank 0's parameters are broadcast at construction and every backward leaves the cross-rank
gradient in place.  Here the exchange is ONE all_reduce(SUM) of the flat gradient bucket
(1.7 MB) on the process group (NCCL over NVLink/NVSwitch on the GPU box); the 1/world_size
is folded into the fused SGD kernel (exact for power-of-two world sizes)."""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn as nn

from .net import Net


class DistributedDataParallel(nn.Module):
    def __init__(self, module: Net, process_group=None):
        super().__init__()
        if not isinstance(module, Net):
            raise TypeError("this DistributedDataParallel wraps pytorch_operator_b200.Net only")
        if not dist.is_initialized():
            raise RuntimeError("init_process_group() first (mnist.py:116)")
        self.module = module
        self.process_group = process_group
        self.world_size = dist.get_world_size(process_group)
        object.__setattr__(module, "_ddp", self)  # plain back-reference, not a registered sub-module (no cycle)
        dist.broadcast(module.flat_params, src=0, group=process_group)  # DDP ctor semantics

    def reduce_gradients(self):
        """Sum the flat gradient bucket over all ranks, in place, on the current stream."""
        dist.all_reduce(self.module.flat_grads, op=dist.ReduceOp.SUM, group=self.process_group)

    def forward(self, x):
        return self.module(x)

    def train_step(self, x, target, optimizer):
        return self.module.train_step(x, target, optimizer)

    def evaluate(self, *a, **k):
        return self.module.evaluate(*a, **k)

    def release_graphs(self):
        self.module.release_graphs()

    @property
    def device(self):
        return self.module.device
