"""Worker for tests/test_gpu_multi.py: one rank of a world_size-2 DDP run on two GPUs (NCCL),
launched with the operator's env contract.  Replays the reference's ws=2 fixture (2 free-running
steps, per-rank batches) and checks losses, averaged gradients and parameters; then runs the
smoke-dist send/recv + all_reduce check.  Exit code 0 = pass."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist

import pytorch_operator_b200 as pkg
from pytorch_operator_b200 import dist_sendrecv
from tests.helpers import RTOL, golden_batch, golden_tensor_check, load_golden


def property_run(mode):
    """Any world size (2/4/8): no reference fixture exists for ws > 2, so check what must hold for every ws --
    the bucket every rank ends up with is the rank-ordered sum of the per-rank gradients (recomputed here from an
    all_gather), all replicas stay bit-identical, and the loss decreases over a few steps."""
    rank, ws = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", 0 if torch.cuda.device_count() == 1 else rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    from oracle import mnist_oracle as mo
    torch.manual_seed(1)
    net = pkg.Net().to(dev)
    model = pkg.DistributedDataParallel(net, fused=mode.startswith("prop-fused"))
    opt = pkg.SGD(model.parameters(), lr=0.01, momentum=0.5)
    losses = []
    for s in range(6):
        x, t = mo.synthetic_batch(64, 5000 + 31 * s + rank)
        local = pkg.Net().to(dev)                      # this rank's own gradient, without any exchange
        local.flat_params.copy_(net.flat_params)
        local.forward_backward(x.to(dev), t.to(dev))
        gathered = [torch.empty_like(local.flat_grads) for _ in range(ws)]
        dist.all_gather(gathered, local.flat_grads)
        want = gathered[0].clone()
        for r in range(1, ws):
            want += gathered[r]                        # rank order, like the fused kernel
        losses.append(float(model.train_step(x.to(dev), t.to(dev), opt)))
        if model.fused:
            assert torch.equal(net.flat_grads, want), f"step {s}: fused sum differs from the rank-ordered sum"
        else:
            torch.testing.assert_close(net.flat_grads, want, rtol=1e-5, atol=1e-7)   # NCCL's own summation order
        ref = net.flat_params.clone()
        dist.broadcast(ref, src=0)
        assert torch.equal(ref, net.flat_params), f"step {s}: replicas diverged"
    lt = torch.tensor([losses[0], losses[-1]], device=dev)
    dist.all_reduce(lt)
    assert float(lt[1]) < float(lt[0]), losses
    rows = dist_sendrecv.allreduce_check(sizes=(4, 431296, 1 << 22), iters=5)
    assert all(r["bit_exact"] for r in rows)
    torch.manual_seed(0)
    dist_sendrecv.run()
    model.release_graphs()
    dist.barrier()
    dist.destroy_process_group()
    print(f"rank {rank} ok ({mode}, ws={ws})", flush=True)


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "eager"
    if mode.startswith("prop"):
        return property_run(mode)
    rank, ws = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    # NCCL modes: the launcher pins one GPU per process (CUDA_VISIBLE_DEVICES), like one pod -> device 0.
    # fused modes: all GPUs stay visible (peer mapping needs distinct device ordinals) -> device = rank.
    dev = torch.device("cuda", 0 if torch.cuda.device_count() == 1 else rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    gold = load_golden("train_ws2_b64")
    assert ws == int(gold["meta/world_size"][0])
    torch.manual_seed(1 + 7 * rank)  # ranks start different; the DDP constructor must broadcast rank 0
    net = pkg.Net().to(dev)
    if rank == 0:
        torch.manual_seed(1)
        net.load_named({k: v for k, v in pkg.Net().state_dict().items()})
    model = pkg.DistributedDataParallel(net, fused=mode.startswith("fused"))
    assert model.fused == mode.startswith("fused"), model.fused_error
    opt = pkg.SGD(model.parameters(), lr=0.01, momentum=0.5)
    graph = None
    for s in range(int(gold["meta/nsteps"][0])):
        x, t = golden_batch(gold, 64, s, rank)
        if mode.endswith("graph") and s > 0:
            graph = graph or pkg.StepGraph(model, opt, 64)
            graph.x.copy_(x)
            graph.target.copy_(t)
            loss = graph.replay()
        else:
            loss = model.train_step(x.to(dev), t.to(dev), opt)
        want = float(gold[f"step{s}/loss_rank{rank}"][0])
        assert abs(float(loss) - want) <= 4 * RTOL * 2.4, (rank, s, float(loss), want)
        if rank == 0:
            # the bucket holds the SUM over ranks; the reference's p.grad is the mean (1/ws is in the SGD kernel)
            for n, v in net.named_views(net.flat_grads).items():
                golden_tensor_check(gold, f"step{s}/grad", n, (v * (1.0 / ws)).cpu().numpy(), rtol=4 * RTOL)
            for n, v in net.named_views(net.flat_params).items():
                golden_tensor_check(gold, f"step{s}/param", n, v.cpu().numpy())
    # both ranks hold identical parameters
    mine = net.flat_params.clone()
    dist.broadcast(mine, src=0)
    assert torch.equal(mine, net.flat_params)
    # smoke-dist on the same process group
    torch.manual_seed(0)
    inp, results = dist_sendrecv.run()
    if rank == 0:
        assert len(results) == ws - 1 and all(torch.equal(r, torch.mul(inp, inp)) for r in results)
    rows = dist_sendrecv.allreduce_check(sizes=(4, 431296, 1 << 20), iters=3)
    assert all(r["bit_exact"] for r in rows)
    model.release_graphs()
    dist.barrier()
    dist.destroy_process_group()
    print(f"rank {rank} ok ({mode})", flush=True)


if __name__ == "__main__":
    main()
