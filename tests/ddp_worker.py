"""Multi-GPU DDP parity checks (one rank's share; NCCL process group already initialised).

Used twice: as the worker script of tests/test_gpu_multi.py (launched with the operator's env contract), and by
bench.py's parity prologue at N>1, which puts the outcome on the driver's record.  Every check raises
AssertionError on a mismatch.

  * ``fixture_run``  -- world_size 2: replays the fixture the unmodified reference produced with
    DistributedDataParallelCPU on gloo (tests/golden/train_ws2_b64.npz; 2 free-running steps, per-rank batches) and
    checks losses, averaged gradients and parameters (mnist.py:135-138,39-43);
  * ``property_run`` -- any world size: no reference fixture exists beyond ws=2, so check what must hold for every
    ws -- the bucket every rank ends up with is the rank-ordered sum of the per-rank gradients (recomputed from an
    all_gather), all replicas stay bit-identical, the loss decreases;
  * ``smoke_dist_run`` -- dist_sendrecv.py:15-34 (send/recv star + square, bit-exact) and the all_reduce
    bit-exactness check.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist

import pytorch_operator_b200 as pkg
from pytorch_operator_b200 import dist_sendrecv
from tests.helpers import RTOL, golden_tensor_check, load_golden


def synthetic_batch(batch_size, seed):
    """SURVEY.md 8(d) synthetic batch (the generator oracle/gen_golden.py and every parity test use)."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(batch_size, 1, 28, 28, generator=g)
    return x, torch.randint(0, 10, (batch_size,), generator=g)


def fixture_run(dev, fused, graph):
    rank, ws = dist.get_rank(), dist.get_world_size()
    gold = load_golden("train_ws2_b64")
    assert ws == int(gold["meta/world_size"][0]), "the reference fixture is a world_size-2 run"
    torch.manual_seed(1 + 7 * rank)  # ranks start different; the DDP constructor must broadcast rank 0
    net = pkg.Net().to(dev)
    if rank == 0:
        torch.manual_seed(1)
        net.load_named({k: v for k, v in pkg.Net().state_dict().items()})
    model = pkg.DistributedDataParallel(net, fused=fused)
    assert model.fused == fused, model.fused_error
    opt = pkg.SGD(model.parameters(), lr=0.01, momentum=0.5)
    g = None
    for s in range(int(gold["meta/nsteps"][0])):
        x, t = synthetic_batch(64, int(gold[f"step{s}/seed_rank{rank}"][0]))
        if graph and s > 0:
            g = g or pkg.StepGraph(model, opt, 64)
            g.x.copy_(x)
            g.target.copy_(t)
            loss = g.replay()
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
    mine = net.flat_params.clone()   # both ranks hold identical parameters
    dist.broadcast(mine, src=0)
    assert torch.equal(mine, net.flat_params), "replicas diverged"
    model.release_graphs()
    model.close()


def property_run(dev, fused, graph=False, steps=6, ring=False):
    rank, ws = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(1)
    net = pkg.Net().to(dev)
    model = pkg.DistributedDataParallel(net, fused=fused)
    assert model.fused == fused, model.fused_error
    opt = pkg.SGD(model.parameters(), lr=0.01, momentum=0.5)
    local = pkg.Net().to(dev)                          # this rank's own gradient, without any exchange
    losses, g = [], None
    for s in range(steps):
        # one batch per rank, repeated: the gradients still change every step (the parameters move), and the
        # loss on it must go down -- on fresh random-label batches it would only wander around ln(10)
        x, t = synthetic_batch(64, 5000 + rank)
        x, t = x.to(dev), t.to(dev)
        local.flat_params.copy_(net.flat_params)
        local.forward_backward(x, t)
        gathered = [torch.empty_like(local.flat_grads) for _ in range(ws)]
        dist.all_gather(gathered, local.flat_grads)
        want = gathered[0].clone()
        for r in range(1, ws):
            want += gathered[r]                        # rank order, like the fused kernels
        if graph and s > 0:
            if g is None:   # ring=True: the replays read this rank's batch from a one-slot resident ring instead
                in_ring = pkg.InputRing(64, x=x[None].clone(), target=t[None].clone()) if ring else None
                g = pkg.StepGraph(model, opt, 64, ring=in_ring)
            if not ring:
                g.x.copy_(x)
                g.target.copy_(t)
            losses.append(float(g.replay()))
        else:
            losses.append(float(model.train_step(x, t, opt)))
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
    model.release_graphs()
    model.close()


def soak_run(dev, steps=400):
    """world_size 2 only: `steps` CUDA-graph steps on the fused path (resident input ring, four steps per graph)
    against the same steps on the NCCL path (one step per graph, batches copied in), bit for bit.  With two ranks every summation order gives the same bits (a + b), the SGD kernels round identically, so the
    two trajectories must stay IDENTICAL -- any race in the fused exchange (a flag seen before its data, a staging
    bucket or receive slot reused too early, a stale parameter slice) shows up as a differing bit within a few steps."""
    rank, ws = dist.get_rank(), dist.get_world_size()
    assert ws == 2
    batches = [synthetic_batch(64, 9000 + 2 * i + rank) for i in range(8)]
    batches = [(x.to(dev), t.to(dev)) for x, t in batches]
    finals = []
    for fused in (True, False):
        torch.manual_seed(1)
        net = pkg.Net().to(dev)
        model = pkg.DistributedDataParallel(net, fused=fused)
        assert model.fused == fused, model.fused_error
        opt = pkg.SGD(model.parameters(), lr=0.01, momentum=0.5)
        model.train_step(*batches[0], opt)
        losses = []
        if fused:
            # the fused run reads its batches from a resident ring, four steps per graph launch (the staging parity
            # alternates inside the graph); the last three steps go out one by one
            order = [batches[(j + 1) % 8] for j in range(8)]
            ring = pkg.InputRing(64, x=torch.stack([b[0] for b in order]), target=torch.stack([b[1] for b in order]))
            g4 = pkg.StepGraph(model, opt, 64, ring=ring, unroll=4)
            g1 = pkg.StepGraph(model, opt, 64, x=g4.x, target=g4.target, ring=ring)
            for r in range(1, (steps - 1) // 4 + 1):
                g4.replay()
                if (4 * r) % 100 == 0:
                    losses.append(float(g4.loss))
            for _ in range((steps - 1) % 4):
                g1.replay()
        else:
            g = pkg.StepGraph(model, opt, 64)
            for i in range(1, steps):
                g.x.copy_(batches[i % 8][0], non_blocking=True)
                g.target.copy_(batches[i % 8][1], non_blocking=True)
                g.replay()
                if i % 100 == 0:
                    losses.append(float(g.loss))
        torch.cuda.synchronize()
        finals.append((net.flat_params.clone(), net.flat_grads.clone(), losses))
        model.release_graphs()
        model.close()
    assert finals[0][2] == finals[1][2], ("losses differ", finals[0][2], finals[1][2])
    assert torch.equal(finals[0][1], finals[1][1]), "last gradient sum differs between the fused and the NCCL path"
    assert torch.equal(finals[0][0], finals[1][0]), "parameters differ between the fused and the NCCL path"
    assert finals[0][2][-1] < finals[0][2][0], finals[0][2]


def smoke_dist_run(sizes=(4, 431296, 1 << 22), iters=5):
    rank, ws = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(0)
    inp, results = dist_sendrecv.run()
    if rank == 0:
        assert len(results) == ws - 1 and all(torch.equal(r, torch.mul(inp, inp)) for r in results)
    rows = dist_sendrecv.allreduce_check(sizes=sizes, iters=iters)
    assert all(r["bit_exact"] for r in rows)
    return rows


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "eager"
    rank = int(os.environ["RANK"])
    # pinned (CUDA_VISIBLE_DEVICES=<rank>, the one-GPU-per-pod view): device 0; all GPUs visible: device = rank
    dev = torch.device("cuda", 0 if torch.cuda.device_count() == 1 else rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    if mode == "soak":
        soak_run(dev)
    elif mode.startswith("prop"):
        property_run(dev, fused=mode.startswith("prop-fused"), graph=mode.endswith("graph") or mode.endswith("ring"),
                     ring=mode.endswith("ring"))
    else:
        fixture_run(dev, fused=mode.startswith("fused"), graph=mode.endswith("graph"))
    smoke_dist_run()
    dist.barrier()
    dist.destroy_process_group()
    print(f"rank {rank} ok ({mode}, ws={os.environ['WORLD_SIZE']})", flush=True)


if __name__ == "__main__":
    main()
