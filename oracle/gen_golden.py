"""Generate tests/golden/*.npz by running the UNMODIFIED reference payloads.

TEST INFRASTRUCTURE ONLY; runs in the build container (needs /root/reference,
which does not exist on the GPU box -- that is why the outputs are committed).

    python oracle/gen_golden.py            # rewrites tests/golden/

What is executed is the reference's own code: ``Net``, ``train`` and ``test`` from
/root/reference/examples/mnist/mnist.py and ``run`` from
/root/reference/examples/smoke-dist/dist_sendrecv.py, imported as they are.
Two obstacles are worked around without touching them:
``tensorboardX`` is not installed, so a stub module whose SummaryWriter records
the scalars is placed in ``sys.modules`` before the import; the MNIST download
needs a network, so ``main()`` is never called and ``train()``/``test()`` get a
list-like loader of seeded synthetic batches instead.

fc1.weight-shaped tensors (400,000 values) are stored as a stride-16 sample plus
float64 sum / L2 norm to keep the fixtures small; every other tensor is stored
whole.  Produced with the torch version recorded in each file.
"""
import importlib.util
import logging
import os
import sys
import types

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, ROOT)

FC1_STRIDE = 16


class _RecordingWriter:
    def __init__(self, *a, **k):
        self.scalars = []

    def add_scalar(self, tag, value, step):
        self.scalars.append((tag, float(value), int(step)))


def _import_reference(world_size: int):
    os.environ["WORLD_SIZE"] = str(world_size)  # mnist.py:14 reads it at import time
    stub = types.ModuleType("tensorboardX")
    stub.SummaryWriter = _RecordingWriter
    sys.modules["tensorboardX"] = stub
    spec = importlib.util.spec_from_file_location("ref_mnist", os.path.join(REF, "examples/mnist/mnist.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class _Loader(list):
    """list of (data, target) with the ``.dataset`` length train()/test() print."""

    def __init__(self, batches):
        super().__init__(batches)
        self.dataset = range(sum(len(b[0]) for b in batches))


def batch_seed(step: int, rank: int) -> int:
    return 1234 + 97 * step + rank


MIN_POOL_GAP = 5e-7


def tie_safe_batch(named_params, batch_size, step, rank):
    """First seed batch_seed(step, rank) + 1000*k whose batch has no max-pool near-tie under the
    current parameters (mnist_oracle.pool_tie_gap): where two window candidates differ by less
    than fp32 summation noise, the arg-max the reference routes the gradient to is an accident
    of its summation order and no other implementation can be expected to reproduce it."""
    from oracle.mnist_oracle import pool_tie_gap, synthetic_batch
    params = {k.replace("module.", ""): v.detach() for k, v in named_params}
    for k in range(2000):
        seed = batch_seed(step, rank) + 1000 * k
        x, t = synthetic_batch(batch_size, seed)
        if pool_tie_gap(params, x) >= MIN_POOL_GAP:
            return seed, x, t
    raise RuntimeError("no tie-safe seed found")


def _pack(prefix, named, out):
    for k, v in named.items():
        a = v.detach().cpu().numpy().astype(np.float32)
        if k.endswith("fc1.weight"):
            flat = a.reshape(-1)
            out[f"{prefix}/{k}/sample"] = flat[::FC1_STRIDE].copy()
            out[f"{prefix}/{k}/sum_l2"] = np.array([flat.astype(np.float64).sum(),
                                                    np.sqrt((flat.astype(np.float64) ** 2).sum())])
        else:
            out[f"{prefix}/{k}"] = a


def _args(log_interval=10):
    return types.SimpleNamespace(log_interval=log_interval)


def _train_case(rank, world_size, batch_size, nsteps, port, result_path):
    from oracle.mnist_oracle import synthetic_batch
    ref = _import_reference(world_size)
    torch.manual_seed(1)  # mnist.py:110 with the default --seed
    if world_size > 1:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                          RANK=str(rank), WORLD_SIZE=str(world_size))
        dist.init_process_group(backend="gloo")  # mnist.py:116, default backend
    model = ref.Net()
    out = {}
    if rank == 0:
        _pack("init", dict(model.named_parameters()), out)
    if ref.is_distributed():
        model = torch.nn.parallel.DistributedDataParallelCPU(model)  # mnist.py:136-138
    opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.5)  # mnist.py:140 defaults
    writer = _RecordingWriter()
    for s in range(nsteps):
        seed, x, t = tie_safe_batch(model.named_parameters(), batch_size, s, rank)
        out[f"step{s}/seed_rank{rank}"] = np.array([seed])
        loader = _Loader([(x, t)])
        ref.train(_args(), model, torch.device("cpu"), loader, opt, 1, writer)  # one step
        names = {k.replace("module.", ""): v for k, v in model.named_parameters()}
        out[f"step{s}/loss_rank{rank}"] = np.array([writer.scalars[-1][1]], dtype=np.float64)
        if rank == 0:
            _pack(f"step{s}/grad", {k: v.grad for k, v in names.items()}, out)
            _pack(f"step{s}/param", names, out)
    np.savez(result_path + f".rank{rank}.npz", **out)
    if world_size > 1:
        dist.barrier()
        dist.destroy_process_group()


def gen_train(name, world_size, batch_size, nsteps, port):
    tmp = os.path.join("/tmp", f"golden_{name}")
    if world_size == 1:
        _train_case(0, 1, batch_size, nsteps, port, tmp)
    else:
        mp.spawn(_train_case, args=(world_size, batch_size, nsteps, port, tmp), nprocs=world_size, join=True)
    merged = {}
    for r in range(world_size):
        with np.load(tmp + f".rank{r}.npz") as z:
            merged.update({k: z[k] for k in z.files})
    merged["meta/world_size"] = np.array([world_size])
    merged["meta/batch_size"] = np.array([batch_size])
    merged["meta/nsteps"] = np.array([nsteps])
    merged["meta/fc1_stride"] = np.array([FC1_STRIDE])
    merged["meta/torch"] = np.array(torch.__version__)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **merged)
    print("wrote", name, {k: v for k, v in merged.items() if "loss" in k})


def gen_eval():
    """reference test() (mnist.py:51-65) over two synthetic batches of 1000."""
    from oracle.mnist_oracle import synthetic_batch
    ref = _import_reference(1)
    torch.manual_seed(1)
    model = ref.Net()
    batches = [synthetic_batch(1000, 4321), synthetic_batch(1000, 4322)]
    writer = _RecordingWriter()
    ref.test(types.SimpleNamespace(), model, torch.device("cpu"), _Loader(batches), writer, 1)
    model.eval()
    with torch.no_grad():
        logp = model(batches[0][0])
    out = {"accuracy": np.array([writer.scalars[-1][1]]), "logp_batch0": logp.numpy(),
           "meta/torch": np.array(torch.__version__)}
    np.savez_compressed(os.path.join(OUT, "eval_b1000.npz"), **out)
    print("wrote eval", out["accuracy"])


class _Capture(logging.Handler):
    def __init__(self):
        super().__init__()
        self.got = []

    def emit(self, record):
        if record.args and len(record.args) == 2 and torch.is_tensor(record.args[1]):
            self.got.append((int(record.args[0]), record.args[1].clone()))


def _smoke_case(rank, world_size, port, result_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world_size))
    spec = importlib.util.spec_from_file_location(
        "ref_sendrecv", os.path.join(REF, "examples/smoke-dist/dist_sendrecv.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    cap = _Capture()
    logging.getLogger().addHandler(cap)
    logging.getLogger().setLevel(logging.INFO)
    torch.manual_seed(0)  # dist_sendrecv.py:19 is unseeded; seeded here so it is reproducible
    mod.init_processes(mod.run)  # dist_sendrecv.py:36-39, gloo
    if rank == 0:
        torch.manual_seed(0)
        inp = torch.randn(2, 2)
        out = {"inp": inp.numpy(), "meta/world_size": np.array([world_size]),
               "meta/torch": np.array(torch.__version__)}
        for w, t in cap.got:
            out[f"result_from_worker{w}"] = t.numpy()
        np.savez(result_path, **out)
    dist.destroy_process_group()


def gen_smoke(world_size=3, port=29631):
    path = os.path.join(OUT, "smoke_sendrecv_ws3.npz")
    mp.spawn(_smoke_case, args=(world_size, port, path), nprocs=world_size, join=True)
    with np.load(path) as z:
        print("wrote smoke", z.files)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(4)
    gen_train("train_ws1_b64", 1, 64, 3, 29611)
    gen_train("train_ws1_b5", 1, 5, 1, 29612)
    gen_train("train_ws2_b64", 2, 64, 2, 29613)
    gen_eval()
    gen_smoke()
