"""CPU-only checks (pytest -m "not gpu"): the C-ABI library loads and exports every symbol the
header declares, host-side argument validation, the operator env contract, the payload CLI,
and the world_size-2 host logic of the flat-bucket DDP on gloo.  No kernel is launched here."""
import ctypes
import json
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

import pytorch_operator_b200 as pkg
from pytorch_operator_b200 import _cabi, launcher

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "b200mnist.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200mnist_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_cabi.build())
    names = _header_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b200mnist.h but not exported"
    assert set(names) == set(_cabi.SIGNATURES), "ctypes signature table out of sync with the header"


def test_flat_layout_is_aligned_and_complete():
    L = _cabi.lib()
    assert L.b200mnist_flat_elems() == 431296
    end = 0
    total = 0
    for i, shape in enumerate(pkg.net.SHAPES):
        off, cnt = L.b200mnist_tensor_offset(i), L.b200mnist_tensor_elems(i)
        assert cnt == int(np.prod(shape)) and off % 64 == 0 and off >= end
        end = off + cnt
        total += cnt
    assert total == 431080 and end <= 431296
    assert L.b200mnist_tensor_offset(8) == -1 and L.b200mnist_tensor_elems(-1) == -1


def test_argument_validation_without_a_gpu():
    L = _cabi.lib()
    assert L.b200mnist_workspace_bytes(0) == -1 and L.b200mnist_workspace_bytes(64) > 64 * 2880 * 4
    assert L.b200mnist_workspace_bytes(64) % 256 == 0
    assert L.b200mnist_sgd_momentum(0, 0, 0, 10, 0.1, 0.5, 1.0, 1, 0) == -10001
    assert "null" in _cabi.last_error()
    assert L.b200mnist_sgd_momentum(16, 16, 16, -1, 0.1, 0.5, 1.0, 1, 0) == -10002
    assert L.b200mnist_sgd_momentum(16, 20, 16, 4, 0.1, 0.5, 1.0, 1, 0) == -10003  # misaligned
    assert L.b200mnist_sgd_momentum(16, 16, 16, 0, 0.1, 0.5, 1.0, 1, 0) == 0       # empty input is a no-op
    assert L.b200mnist_square(0, 0, 0, 0) == 0
    assert L.b200mnist_conv1_relu_pool_fwd(16, 16, 16, 16, 16, 0, 0) == -10002
    assert L.b200mnist_forward_backward(16, 16, 16, 16, 70000, 256, 1 << 40, 16, 0, 0) == -10002


def test_net_mirrors_reference_module_without_a_gpu():
    torch.manual_seed(1)
    net = pkg.Net()
    ref = torch.nn.Sequential()  # reference layer order, same seed
    torch.manual_seed(1)
    layers = [torch.nn.Conv2d(1, 20, 5, 1), torch.nn.Conv2d(20, 50, 5, 1), torch.nn.Linear(800, 500),
              torch.nn.Linear(500, 10)]
    want = {}
    for name, l in zip(("conv1", "conv2", "fc1", "fc2"), layers):
        want[name + ".weight"], want[name + ".bias"] = l.weight, l.bias
    sd = net.state_dict()
    assert list(sd) == pkg.NAMES
    for k in pkg.NAMES:
        assert torch.equal(sd[k], want[k].detach())
    assert sum(p.numel() for p in net.parameters()) == 431080
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        net(torch.zeros(2, 1, 28, 28))
    opt = pkg.SGD(net.parameters(), lr=0.01, momentum=0.5)
    assert opt.model is net and opt.first_step
    with pytest.raises(ValueError):
        pkg.SGD([torch.nn.Parameter(torch.zeros(3))], lr=0.1)


# ---- operator env contract: golden values of TestClusterSpec, pod_test.go:108-150 ----
@pytest.mark.parametrize("workers,rtype,index,expect", [
    (0, "Master", 0, {"WORLD_SIZE": "1", "MASTER_PORT": "23456", "RANK": "0", "MASTER_ADDR": "localhost"}),
    (1, "Master", 0, {"WORLD_SIZE": "2", "MASTER_PORT": "23456", "RANK": "0", "MASTER_ADDR": "localhost"}),
    (1, "Worker", 0, {"WORLD_SIZE": "2", "MASTER_PORT": "23456", "RANK": "1", "MASTER_ADDR": "test-pytorchjob-master-0"}),
    (2, "Master", 0, {"WORLD_SIZE": "3", "MASTER_PORT": "23456", "RANK": "0", "MASTER_ADDR": "localhost"}),
    (2, "Worker", 0, {"WORLD_SIZE": "3", "MASTER_PORT": "23456", "RANK": "1", "MASTER_ADDR": "test-pytorchjob-master-0"}),
    (2, "Worker", 1, {"WORLD_SIZE": "3", "MASTER_PORT": "23456", "RANK": "2", "MASTER_ADDR": "test-pytorchjob-master-0"}),
])
def test_cluster_spec_matches_operator_golden(workers, rtype, index, expect):
    env = launcher.cluster_spec("test-pytorchjob", workers, rtype, index)
    for k, v in expect.items():
        assert env[k] == v
    assert env["PYTHONUNBUFFERED"] == "0"


def test_cluster_spec_rejects_second_master():
    with pytest.raises(ValueError):
        launcher.cluster_spec("j", 1, "Master", 1)


def test_payload_cli_and_helpers():
    env = dict(os.environ, WORLD_SIZE="2", PYTHONPATH=ROOT)
    code = ("import pytorch_operator_b200.mnist as m; print(m.WORLD_SIZE, m.should_distribute(), m.is_distributed());"
            "import inspect; print(list(inspect.signature(m.train).parameters));"
            "print(list(inspect.signature(m.test).parameters))")
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, check=True).stdout
    lines = out.strip().splitlines()
    assert lines[0] == "2 True False"
    assert lines[1] == str(["args", "model", "device", "train_loader", "optimizer", "epoch", "writer"])
    assert lines[2] == str(["args", "model", "device", "test_loader", "writer", "epoch"])
    # no GPU here: the payload must refuse to run rather than fall back to the CPU
    r = subprocess.run([sys.executable, "-m", "pytorch_operator_b200.mnist", "--synthetic", "2"],
                       env=dict(os.environ, PYTHONPATH=ROOT, WORLD_SIZE="1"), capture_output=True, text=True)
    if not torch.cuda.is_available():
        assert r.returncode != 0 and "no CPU fallback" in (r.stderr + r.stdout)
    r = subprocess.run([sys.executable, "-m", "pytorch_operator_b200.mnist", "--help"],
                       env=dict(os.environ, PYTHONPATH=ROOT), capture_output=True, text=True, check=True)
    for flag in ("--batch-size", "--test-batch-size", "--epochs", "--lr", "--momentum", "--no-cuda", "--seed",
                 "--log-interval", "--save-model", "--dir", "--backend"):
        assert flag in r.stdout


def _ddp_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(launcher.cluster_spec("t", world - 1, "master" if rank == 0 else "worker", max(rank - 1, 0), port))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    dist.init_process_group("gloo")
    torch.manual_seed(100 + rank)  # different initial weights per rank: the ctor broadcast must fix that
    net = pkg.Net()
    ddp = pkg.DistributedDataParallel(net)
    opt = pkg.SGD(ddp.parameters(), lr=0.01, momentum=0.5)
    torch.manual_seed(100)
    want = pkg.Net().flat_params
    ok_bcast = torch.equal(net.flat_params, want)
    net.flat_grads.copy_(torch.arange(net.flat_grads.numel(), dtype=torch.float32) % 7 + rank)
    ddp.reduce_gradients()
    expect = (torch.arange(net.flat_grads.numel(), dtype=torch.float32) % 7) * world + sum(range(world))
    ok_sum = torch.equal(net.flat_grads, expect)
    ok_views = torch.equal(net.fc2.bias.grad, expect[431232:431242])
    q.put((rank, ok_bcast, ok_sum, ok_views, opt.effective_grad_scale(), list(ddp.state_dict())[0]))
    dist.barrier()
    dist.destroy_process_group()


def test_flat_bucket_ddp_host_logic_world_size_2_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_ddp_worker, args=(r, 2, 29641, q)) for r in range(2)]
    [p.start() for p in procs]
    res = sorted(q.get(timeout=120) for _ in procs)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    for rank, ok_bcast, ok_sum, ok_views, scale, first_key in res:
        assert ok_bcast and ok_sum and ok_views, (rank, ok_bcast, ok_sum, ok_views)
        assert scale == 0.5 and first_key == "module.conv1.weight"


def test_reference_arm_survives_a_torchrun_environment():
    """bench.py --impl reference at N>1 must not inherit the elastic agent's rendezvous (round 1: the gloo
    children connected to the agent's store and the arm hung for 812 s): run it with the variables torchrun
    exports to a worker, pointing at a port where nobody listens."""
    env = dict(os.environ, PYTHONPATH=ROOT, TORCHELASTIC_USE_AGENT_STORE="True", TORCHELASTIC_RUN_ID="x",
               MASTER_ADDR="127.0.0.1", MASTER_PORT="1", RANK="0", WORLD_SIZE="2", LOCAL_RANK="0",
               GROUP_RANK="0", OMP_NUM_THREADS="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                        "--steps", "3", "--warmup", "3", "--ref-budget", "1"], env=env, capture_output=True, text=True,
                       timeout=240)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["value"] > 0
    assert "2 gloo ranks" in line["cpu_baseline"]["sample"] and "failed" not in line["config"]["how"]
    # every other rank of the torchrun group exits 0 without work and without output
    r1 = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                        env=dict(env, RANK="1", LOCAL_RANK="1"), capture_output=True, text=True, timeout=60)
    assert r1.returncode == 0 and r1.stdout.strip() == ""


@pytest.mark.parametrize("name,module,args", [
    ("pytorch_job_mnist_b200_nccl.yaml", "pytorch_operator_b200.mnist", ["--backend", "nccl"]),
    ("pytorch_job_smoke_dist_b200.yaml", "pytorch_operator_b200.dist_sendrecv", []),
])
def test_pytorchjob_manifests_follow_the_operator_contract(name, module, args):
    """The B200 manifests against what the operator requires of a PyTorchJob: container named "pytorch"
    (pkg/apis/pytorch/v1/constants.go:29, validation.go:58-62), port name "pytorchjob-port" = 23456
    (constants.go:27,31), exactly one Master (manifests/crd.yaml:28-33), restartPolicy OnFailure (constants.go:33),
    one GPU per replica as in examples/mnist/v1/pytorch_job_mnist_nccl.yaml -- and 1 + 7 replicas for one 8-GPU node,
    whose env the launcher reproduces."""
    import yaml
    with open(os.path.join(ROOT, "manifests", name)) as f:
        job = yaml.safe_load(f)
    assert job["apiVersion"] == "kubeflow.org/v1" and job["kind"] == "PyTorchJob"
    specs = job["spec"]["pytorchReplicaSpecs"]
    assert set(specs) == {"Master", "Worker"} and specs["Master"]["replicas"] == 1 and specs["Worker"]["replicas"] == 7
    for rs in specs.values():
        assert rs["restartPolicy"] == "OnFailure"
        (c,) = rs["template"]["spec"]["containers"]
        assert c["name"] == "pytorch" and c["command"] == ["python", "-m", module] and c["args"] == args
        assert c["ports"] == [{"name": "pytorchjob-port", "containerPort": launcher.DEFAULT_PORT}]
        assert c["resources"]["limits"]["nvidia.com/gpu"] == 1
    total = specs["Master"]["replicas"] + specs["Worker"]["replicas"]
    envs = [launcher.cluster_spec(job["metadata"]["name"], 7, "master", 0)] + \
           [launcher.cluster_spec(job["metadata"]["name"], 7, "worker", i) for i in range(7)]
    assert [e["RANK"] for e in envs] == [str(r) for r in range(total)] and {e["WORLD_SIZE"] for e in envs} == {"8"}
    assert envs[1]["MASTER_ADDR"] == job["metadata"]["name"] + "-master-0" and envs[0]["MASTER_ADDR"] == "localhost"


def test_peer_exchange_entry_points_reject_bad_arguments_without_a_gpu():
    """The fused DDP step's C entry points validate before they touch CUDA: bad world sizes, ranks, parities, null or
    misaligned windows come back as argument errors (-10000 - k) with a message, never as a launch."""
    L = _cabi.lib()
    assert L.b200mnist_peer_window_bytes() == 4096 + 4 * 4 * L.b200mnist_flat_elems() + 2 * 8 * 246272
    win = (ctypes.c_void_p * 8)(*([0x10000] * 8))
    mom, x, t, ws, loss = 0x20000, 0x30000, 0x40000, 0x50000, 0x60000

    def step(world=2, rank=0, parity=0, windows=win, momentum=mom, batch=64, ws_bytes=1 << 40):
        return L.b200mnist_peer_train_step(windows, rank, world, parity, momentum, x, t, batch, ws, ws_bytes, 0.01, 0.5, 0,
                                           loss, 0)

    for kw in (dict(world=3), dict(world=1), dict(world=16), dict(rank=2), dict(rank=-1), dict(parity=2)):
        assert step(**kw) == -10002, kw
        assert _cabi.last_error()
    assert step(windows=(ctypes.c_void_p * 8)(0x10000, 0)) == -10003          # a window is missing
    assert step(windows=(ctypes.c_void_p * 8)(0x10000, 0x10010)) == -10003    # not 256-byte aligned
    assert step(momentum=0) == -10001
    assert step(momentum=mom + 4) == -10003
    assert step(batch=0) == -10002
    assert step(ws_bytes=1024) == -10004 and "workspace" in _cabi.last_error()
    assert L.b200mnist_peer_window_probe(win, 9, 0x20000, 0) == -10002
    assert L.b200mnist_peer_window_open(0, ctypes.byref(ctypes.c_void_p())) == -10001
    assert L.b200mnist_peer_window_close(0) == 0 and L.b200mnist_peer_window_destroy(0) == 0
    # the workspace map of the saved activations
    offs = [L.b200mnist_workspace_offset(64, i) for i in range(6)]
    assert offs == sorted(offs) and offs[0] >= 512 and all(o % 256 == 0 for o in offs)
    assert L.b200mnist_workspace_offset(64, 6) == -1 and L.b200mnist_workspace_offset(0, 0) == -1
    assert L.b200mnist_workspace_bytes(4096) > L.b200mnist_workspace_bytes(64) > offs[-1]


def test_ddp_worker_batches_are_the_oracles():
    """tests/ddp_worker.py (also bench.py's parity prologue) draws its batches without importing the oracle: same bits."""
    from oracle import mnist_oracle as mo
    from tests import ddp_worker as dw
    for seed in (0, 1234, 9001):
        a, b = dw.synthetic_batch(64, seed), mo.synthetic_batch(64, seed)
        assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])


def test_bench_clock_summary_of_a_region_shorter_than_the_sampling_period():
    """bench.py: with the driver's --steps 20 the timed region lasts ~1 ms, shorter than nvidia-smi's 50 ms sampling
    period; the clock block then reports the samples of the untimed load phase directly before the region and says so
    (a long region reports its own samples and carries no note); throttle reasons are collected either way."""
    sys.path.insert(0, ROOT)
    import bench
    s = bench.ClockSampler.__new__(bench.ClockSampler)
    row = "1965, 1965, 600.0, Not Active, Not Active, Not Active, {}"
    s.rows = [(100.00 + 0.05 * i, row.format("Active" if i == 6 else "Not Active")) for i in range(10)]   # t = 100.00 .. 100.45
    short = s.under_load(100.452, 100.453)            # no sample inside; the load phase before it has several
    assert short["samples"] >= 5 and short["sm_mhz"] == 1965.0 and "shorter than the 50 ms sampling period" in short["note"]
    assert short["reasons"] == ["sw_power_cap"]
    long = s.under_load(100.0, 100.45)
    assert long["samples"] == 10 and "note" not in long
    s.rows = []
    assert s.under_load(1.0, 2.0) is None               # nvidia-smi missing: no clock block rather than a made-up one


def test_ring_description_and_ring_entry_points_reject_bad_arguments_without_a_gpu():
    """b200mnist_train_step_ring validates the ring description before anything is launched."""
    L = _cabi.lib()
    n = int(L.b200mnist_flat_elems())
    buf = np.zeros(n + 64, dtype=np.float32)
    base = (buf.ctypes.data + 255) // 256 * 256
    ws_bytes = int(L.b200mnist_workspace_bytes(4))
    ws = np.zeros(ws_bytes + 256, dtype=np.uint8)
    wsp = (ws.ctypes.data + 255) // 256 * 256
    call = lambda ring: L.b200mnist_train_step_ring(base, base, base, ring, base, base, 4, wsp, ws_bytes, 0.01, 0.5, 1, base, None)
    assert call(None) < 0 and "ring" in _cabi.last_error()
    ring = _cabi.Ring(base, base, 4 * 784, 4, 0, base, None, None)          # no slots
    assert call(ctypes.byref(ring)) < 0 and "slot" in _cabi.last_error()
    ring = _cabi.Ring(base, base, 4 * 784 - 1, 4, 2, base, None, None)      # slots overlap / misaligned stride
    assert call(ctypes.byref(ring)) < 0 and "aligned" in _cabi.last_error()
    ring = _cabi.Ring(base, base, 4 * 784, 4, 2, None, None, None)          # no cursor
    assert call(ctypes.byref(ring)) < 0 and "null" in _cabi.last_error()
