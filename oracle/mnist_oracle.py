"""CPU oracle for the examples/mnist training step and the examples/smoke-dist check.

TEST INFRASTRUCTURE ONLY.  Nothing under ``pytorch-operator_b200/`` may import this
module; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs do.  It is a restatement, on
``torch`` CPU ops, of what the reference payloads compute -- the arithmetic of
the reference lives in a third-party dependency (PyTorch; image pin
``pytorch/pytorch:1.0-cuda10.0-cudnn7-runtime``, examples/mnist/Dockerfile:1),
so the restatement calls the same library ops the reference calls, in the same
order, but in functional form over an explicit parameter dict.

Parity status: the reference repo holds NO numeric golden vectors for this path
("parity unpinned" by the reference's own tests).  This
oracle is pinned instead against outputs of the reference itself, run in the
build container: ``oracle/gen_golden.py`` imports the unmodified
``/root/reference/examples/mnist/mnist.py`` (``Net``, ``train``) and
``examples/smoke-dist/dist_sendrecv.py`` (``run``) and writes
``tests/golden/*.npz``; ``tests/test_oracle.py`` checks every function here
against those files.
"""
from __future__ import annotations

import math
from collections import OrderedDict

import torch
import torch.nn.functional as F

# Parameter names/shapes in declaration order (examples/mnist/mnist.py:20-23).
PARAM_SHAPES = OrderedDict([
    ("conv1.weight", (20, 1, 5, 5)), ("conv1.bias", (20,)),
    ("conv2.weight", (50, 20, 5, 5)), ("conv2.bias", (50,)),
    ("fc1.weight", (500, 800)), ("fc1.bias", (500,)),
    ("fc2.weight", (10, 500)), ("fc2.bias", (10,)),
])
N_PARAMS = sum(math.prod(s) for s in PARAM_SHAPES.values())  # 431,080


def init_params(seed: int = 1) -> "OrderedDict[str, torch.Tensor]":
    """Initial parameters the reference gets from ``torch.manual_seed(seed); Net()``
    (mnist.py:110,133 with Net.__init__ at :17-23): nn.Conv2d / nn.Linear default
    init is kaiming_uniform_(a=sqrt(5)) on the weight then U(-1/sqrt(fan_in), +) on
    the bias, layer by layer in declaration order.  Restated with the raw init
    calls so the RNG stream is consumed identically."""
    gen_state = torch.random.get_rng_state()
    torch.manual_seed(seed)
    out = OrderedDict()
    names = list(PARAM_SHAPES)
    for wname, bname in zip(names[0::2], names[1::2]):
        w = torch.empty(PARAM_SHAPES[wname])
        torch.nn.init.kaiming_uniform_(w, a=math.sqrt(5))
        fan_in = math.prod(PARAM_SHAPES[wname][1:])
        bound = 1.0 / math.sqrt(fan_in)
        b = torch.empty(PARAM_SHAPES[bname])
        torch.nn.init.uniform_(b, -bound, bound)
        out[wname], out[bname] = w, b
    torch.random.set_rng_state(gen_state)
    return out


def synthetic_batch(batch_size: int, seed: int):
    """Seeded synthetic batch: randn images at normalised-MNIST scale,
    int64 labels in [0,10) drawn from the same generator."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(batch_size, 1, 28, 28, generator=g)
    t = torch.randint(0, 10, (batch_size,), generator=g)
    return x, t


def pool_tie_gap(params, x) -> float:
    """Smallest gap between the two largest values of any live 2x2 max-pool window (both
    pooling stages, mnist.py:27,29), computed in float64.  When two candidates are closer than
    fp32 summation noise, WHICH one max_pool2d's backward routes the gradient to depends on the
    order in which the convolution summed its 25/500 terms -- the reference's own choice is then
    arbitrary -- so gradient parity is only defined up to that choice."""
    pp = {k: v.double() for k, v in params.items()}
    gaps = []
    a = x.double()
    for wn, bn in (("conv1.weight", "conv1.bias"), ("conv2.weight", "conv2.bias")):
        c = F.relu(F.conv2d(a, pp[wn], pp[bn]))
        w = c.unfold(2, 2, 2).unfold(3, 2, 2).reshape(*c.shape[:2], c.shape[2] // 2, c.shape[3] // 2, 4)
        top2 = w.topk(2, dim=-1).values
        live = top2[..., 0] > 0
        gaps.append(float((top2[..., 0] - top2[..., 1])[live].min()) if live.any() else 1.0)
        a = F.max_pool2d(c, 2, 2)
    return min(gaps)


def relu_kink_gap(params, x):
    """(smallest |pre-ReLU value| of any pooled conv2 output, smallest |pre-ReLU value| of any fc1 output), float64:
    how close the batch comes to a ReLU kink of mnist.py:28-31.  A value within rounding of zero is switched on or off
    by the summation order (or, for the 3xTF32 tensor-core kernels, by their ~1e-6 relative error), and that element's
    gradient with it."""
    pp = {k: v.double() for k, v in params.items()}
    a = F.max_pool2d(F.relu(F.conv2d(x.double(), pp["conv1.weight"], pp["conv1.bias"])), 2, 2)
    pre2 = F.max_pool2d(F.conv2d(a, pp["conv2.weight"], pp["conv2.bias"]), 2, 2)
    z1 = F.linear(F.relu(pre2).reshape(-1, 800), pp["fc1.weight"], pp["fc1.bias"])
    return float(pre2.abs().min()), float(z1.abs().min())


def pool_choice_slack(params, x, idx1, idx2) -> float:
    """How far (float64) the window candidates picked by the arg-max maps idx1 [B,20,12,12] /
    idx2 [B,50,4,4] (slot = dy*2+dx) are below the true window maximum, over live windows of
    both pooling stages.  ~1e-7 means the choice is a valid maximum up to fp32 rounding."""
    pp = {k: v.double() for k, v in params.items()}
    a = x.double()
    worst = 0.0
    for wn, bn, idx in (("conv1.weight", "conv1.bias", idx1), ("conv2.weight", "conv2.bias", idx2)):
        c = F.relu(F.conv2d(a, pp[wn], pp[bn]))
        w = c.unfold(2, 2, 2).unfold(3, 2, 2).reshape(*c.shape[:2], c.shape[2] // 2, c.shape[3] // 2, 4)
        mx = w.max(dim=-1).values
        picked = w.gather(-1, torch.as_tensor(idx).long().unsqueeze(-1)).squeeze(-1)
        live = mx > 1e-6
        if live.any():
            worst = max(worst, float((mx - picked)[live].max()))
        a = F.max_pool2d(c, 2, 2)
    return worst


def forward(params, x: torch.Tensor) -> torch.Tensor:
    """Net.forward (mnist.py:25-33): conv5x5+ReLU, 2x2 max-pool, conv5x5+ReLU,
    2x2 max-pool, flatten to 800, fc+ReLU, fc, log_softmax over the 10 classes."""
    a = F.max_pool2d(F.relu(F.conv2d(x, params["conv1.weight"], params["conv1.bias"])), 2, 2)
    a = F.max_pool2d(F.relu(F.conv2d(a, params["conv2.weight"], params["conv2.bias"])), 2, 2)
    a = a.reshape(-1, 800)
    a = F.relu(F.linear(a, params["fc1.weight"], params["fc1.bias"]))
    a = F.linear(a, params["fc2.weight"], params["fc2.bias"])
    return F.log_softmax(a, dim=1)


def loss_and_grads(params, x, target):
    """One forward + ``F.nll_loss`` (mean) + backward (mnist.py:40-42).  Returns
    (loss, logp, grads) with grads keyed like params; params are not modified."""
    leaves = OrderedDict((k, v.detach().clone().requires_grad_(True)) for k, v in params.items())
    logp = forward(leaves, x)
    loss = F.nll_loss(logp, target)
    gl = torch.autograd.grad(loss, list(leaves.values()))
    return loss.detach(), logp.detach(), OrderedDict(zip(leaves.keys(), gl))


def sgd_momentum_(params, grads, bufs, lr=0.01, momentum=0.5):
    """optim.SGD(lr, momentum) step with dampening 0, no nesterov, no weight decay
    (mnist.py:140,43): first step ``buf = g``; afterwards ``buf = mu*buf + g``;
    ``p = p - lr*buf``.  ``bufs`` is a dict that is filled on the first call."""
    for k, p in params.items():
        g = grads[k]
        if k not in bufs:
            bufs[k] = g.clone()
        else:
            bufs[k].mul_(momentum).add_(g)
        p.add_(bufs[k], alpha=-lr)


def average_grads(per_rank_grads):
    """What DDP leaves in ``p.grad`` on every rank (mnist.py:135-138): the sum of the
    ranks' gradients divided by world_size."""
    ws = len(per_rank_grads)
    out = OrderedDict()
    for k in per_rank_grads[0]:
        s = per_rank_grads[0][k].clone()
        for r in range(1, ws):
            s = s + per_rank_grads[r][k]
        out[k] = s / ws
    return out


def train_steps(params, batches_per_rank, lr=0.01, momentum=0.5):
    """Free-running restatement of ``train()`` (mnist.py:35-43) for world_size =
    len(batches_per_rank) replicas that start identical and see per-rank batches.
    ``batches_per_rank[r]`` is a list of (x, target).  Returns per-step
    dicts {losses: [per-rank], grads: averaged grads} and leaves the final
    parameters in ``params`` / momentum in the returned ``bufs``."""
    bufs, trace = {}, []
    nsteps = len(batches_per_rank[0])
    for s in range(nsteps):
        losses, grads = [], []
        for rb in batches_per_rank:
            l, _, g = loss_and_grads(params, *rb[s])
            losses.append(float(l))
            grads.append(g)
        avg = average_grads(grads) if len(grads) > 1 else grads[0]
        sgd_momentum_(params, avg, bufs, lr, momentum)
        trace.append({"losses": losses, "grads": avg})
    return trace, bufs


def evaluate(params, x, target):
    """One batch of ``test()`` (mnist.py:56-61): summed NLL and number of correct
    argmax predictions."""
    logp = forward(params, x)
    loss_sum = float(F.nll_loss(logp, target, reduction="sum"))
    pred = logp.max(1, keepdim=True)[1]
    correct = int(pred.eq(target.view_as(pred)).sum())
    return loss_sum, correct, logp


def smoke_sendrecv(inp: torch.Tensor, world_size: int):
    """What rank 0 receives from each worker in dist_sendrecv.py:15-34: every worker
    returns ``torch.mul(inp, inp)`` of the tensor rank 0 sent it."""
    return [torch.mul(inp, inp) for _ in range(1, world_size)]


class TorchCpuTrainer:
    """Timed CPU baseline for bench.py: the same op sequence as the reference's
    ``train()`` body, on nn.Module-free functional torch CPU ops with autograd and
    ``torch.optim.SGD``, so the library kernels exercised (mkldnn conv, addmm, ...)
    are the ones the reference exercises on a CPU pod."""

    def __init__(self, seed=1, lr=0.01, momentum=0.5):
        self.params = OrderedDict((k, v.requires_grad_(True)) for k, v in init_params(seed).items())
        self.opt = torch.optim.SGD(list(self.params.values()), lr=lr, momentum=momentum)

    def step(self, x, target) -> torch.Tensor:
        self.opt.zero_grad()
        loss = F.nll_loss(forward(self.params, x), target)
        loss.backward()
        self.opt.step()
        return loss.detach()


class OracleNet(torch.nn.Module):
    """Module form of the same network (layer names as in mnist.py:20-23) for the gloo/DDP CPU baseline."""

    def __init__(self):
        super().__init__()
        self.conv1 = torch.nn.Conv2d(1, 20, 5, 1)
        self.conv2 = torch.nn.Conv2d(20, 50, 5, 1)
        self.fc1 = torch.nn.Linear(800, 500)
        self.fc2 = torch.nn.Linear(500, 10)

    def forward(self, x):
        return forward(dict(self.named_parameters()), x)


def ddp_cpu_worker(rank, world, port, batch, steps, warmup, budget_s, threads, result_path):
    """One rank of the reference's CPU/gloo configuration (BASELINE.json configs[0]; mnist.py:116,136-140):
    DistributedDataParallel on gloo, optim.SGD(lr=.01, momentum=.5), per-rank synthetic batches.  Rank 0
    writes {value: global images/s, ...} as JSON to result_path.

    Rendezvous is an explicit ``tcp://127.0.0.1:<port>`` store hosted by rank 0 of THIS group: the caller may
    itself be a torchrun worker, whose TORCHELASTIC_* / MASTER_* variables would otherwise point ``env://`` at
    the elastic agent's store, where nobody serves this group (the round-1 hang)."""
    import datetime
    import json
    import time

    import torch.distributed as dist
    torch.set_num_threads(max(1, threads))
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            timeout=datetime.timedelta(seconds=60))
    torch.manual_seed(1)
    model = torch.nn.parallel.DistributedDataParallel(OracleNet())
    opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.5)
    ring = [synthetic_batch(batch, 1234 + rank + 97 * i) for i in range(16)]

    def step(i):
        x, t = ring[i % 16]
        opt.zero_grad()
        loss = F.nll_loss(model(x), t)
        loss.backward()
        opt.step()
        return loss

    t0 = time.perf_counter()
    for i in range(max(warmup, 3)):
        step(i)
    per = torch.tensor([(time.perf_counter() - t0) / max(warmup, 3)])
    dist.all_reduce(per, op=dist.ReduceOp.MAX)
    n = int(max(min(steps, 200), min(steps, budget_s / float(per))))  # the same count on every rank; at least 200
    dist.barrier()
    t0 = time.perf_counter()
    for i in range(n):
        step(i)
    dist.barrier()
    dt = time.perf_counter() - t0
    if rank == 0:
        with open(result_path, "w") as f:
            json.dump({"value": world * batch * n / dt, "steps": n, "ms_per_step": 1e3 * dt / n, "seconds": dt,
                       "threads_per_rank": max(1, threads), "world": world}, f)
    dist.destroy_process_group()


# environment a torchrun / launcher parent leaves behind that must not leak into the CPU baseline's own group
_SCRUB_PREFIXES = ("TORCHELASTIC_", "TORCH_NCCL_", "NCCL_", "PET_")
_SCRUB_NAMES = ("MASTER_ADDR", "MASTER_PORT", "RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE", "GROUP_RANK",
                "GROUP_WORLD_SIZE", "ROLE_RANK", "ROLE_WORLD_SIZE", "ROLE_NAME", "OMP_NUM_THREADS",
                "CUDA_VISIBLE_DEVICES")


def clean_env(env=None):
    """A copy of ``env`` (default os.environ) without any rendezvous / elastic-agent variable."""
    import os
    env = dict(os.environ if env is None else env)
    for k in list(env):
        if k in _SCRUB_NAMES or k.startswith(_SCRUB_PREFIXES):
            del env[k]
    return env


def run_ddp_cpu(world, batch, steps, warmup, budget_s, threads, hard_timeout_s):
    """Start ``world`` fresh interpreter processes (clean environment, own tcp store on a free port) running
    ``ddp_cpu_worker`` and return rank 0's result dict.  Every child is killed when ``hard_timeout_s`` expires:
    this arm can cost the caller at most that long, whatever goes wrong inside it."""
    import json
    import os
    import socket
    import subprocess
    import sys
    import tempfile
    import time
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    path = os.path.join(tempfile.mkdtemp(), "ddp_cpu.json")
    env = clean_env()
    env["PYTHONPATH"] = root + os.pathsep + env.get("PYTHONPATH", "")
    env["CUDA_VISIBLE_DEVICES"] = ""  # the CPU arm must not touch (or initialise) a GPU
    procs = [subprocess.Popen([sys.executable, "-m", "oracle.mnist_oracle", "ddp-worker", str(r), str(world), str(port),
                               str(batch), str(steps), str(warmup), str(budget_s), str(threads), path],
                              env=env, cwd=root) for r in range(world)]
    deadline = time.monotonic() + hard_timeout_s
    rcs = []
    for p in procs:
        try:
            rcs.append(p.wait(timeout=max(0.1, deadline - time.monotonic())))
        except subprocess.TimeoutExpired:
            for q in procs:
                if q.poll() is None:
                    q.kill()
            raise TimeoutError(f"gloo DDP CPU baseline did not finish within {hard_timeout_s:.0f} s") from None
    if any(rcs):
        raise RuntimeError(f"gloo DDP CPU baseline: worker exit codes {rcs}")
    with open(path) as f:
        return json.load(f)


if __name__ == "__main__":
    import sys
    if len(sys.argv) >= 11 and sys.argv[1] == "ddp-worker":
        a = sys.argv[2:]
        ddp_cpu_worker(int(a[0]), int(a[1]), int(a[2]), int(a[3]), int(a[4]), int(a[5]), float(a[6]), int(a[7]), a[8])
    else:
        raise SystemExit("usage: python -m oracle.mnist_oracle ddp-worker RANK WORLD PORT BATCH STEPS WARMUP BUDGET_S "
                         "THREADS RESULT_PATH")
