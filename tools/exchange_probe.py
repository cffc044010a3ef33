#!/usr/bin/env python
"""Steady-state step time and exchange telemetry of the fused DDP step for a few launch variants (run under
torchrun on N GPUs):  python -m torch.distributed.run --nproc-per-node N tools/exchange_probe.py [batch]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

import pytorch_operator_b200 as pkg

rank, lr = int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", os.environ["RANK"]))
dev = torch.device("cuda", lr % torch.cuda.device_count())
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
torch.manual_seed(1)
net = pkg.Net().to(dev)
model = pkg.DistributedDataParallel(net, fused=True)
opt = pkg.SGD(model.parameters(), lr=0.01, momentum=0.5)
g = torch.Generator(device=dev).manual_seed(7 + rank)
x = torch.randn(B, 1, 28, 28, device=dev, generator=g)
t = torch.randint(0, 10, (B,), device=dev, generator=g)
model.train_step(x, t, opt)
rows = []
for name in ("run1", "run2", "run3"):
    graph = pkg.StepGraph(model, opt, B, x=x, target=t)
    for _ in range(200):
        graph.replay()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3000):
        graph.replay()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / 3000], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    tel = model.exchange_telemetry()
    graph.release()
    dist.barrier()
    if rank == 0:
        rows.append({"variant": name, "us_per_step": round(float(ms) * 1e3, 2), "telemetry_rank0": tel})
        print(json.dumps(rows[-1]), flush=True)
model.close()
dist.barrier()
dist.destroy_process_group()
