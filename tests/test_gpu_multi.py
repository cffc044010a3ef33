"""Multi-GPU parity (BASELINE configs 3 and 4): flat-bucket DDP -- NCCL all_reduce path and the fused peer-window
exchange kernels -- against the fixture the reference produced with DistributedDataParallelCPU on gloo (ws=2), the
rank-ordered-sum / identical-replica properties at ws=2/4/8, the smoke-dist send/recv + all_reduce bit-exactness
check, and the smoke-dist payload's own main() under the launcher.  Each case skips when the box has fewer GPUs than
it needs (the driver's test box has one; bench.py's parity prologue records the same checks at N>1)."""
import os

import pytest
import torch

from pytorch_operator_b200 import launcher

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "ddp_worker.py")
MODES = ["eager", "graph", "fused", "fused-graph"]
PROPS = ["prop-nccl", "prop-fused", "prop-fused-graph", "prop-fused-ring"]


def _need(n):
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs")


@pytest.mark.parametrize("mode", MODES)
def test_ddp_two_gpus_matches_reference_fixture(mode):
    """One process per GPU, each pinned to its GPU with CUDA_VISIBLE_DEVICES like a pod with nvidia.com/gpu: 1 --
    the fused exchange maps its peers through CUDA IPC and must work in that view."""
    _need(2)
    rc = launcher.launch([WORKER, mode], workers=1, port=23456 + MODES.index(mode), timeout=180)
    assert rc == 0


def test_ddp_two_gpus_fused_with_all_gpus_visible():
    """The same fused run with every GPU visible to every process (device = rank), as under torchrun."""
    _need(2)
    rc = launcher.launch([WORKER, "fused-graph"], workers=1, port=23466, timeout=180, gpus=False)
    assert rc == 0


def test_ddp_two_gpus_fused_trajectory_is_bit_identical_to_nccl():
    """400 CUDA-graph steps on the fused exchange vs the same steps on NCCL all_reduce + SGD: identical bits (a race
    detector for the flag-free exchange; see ddp_worker.soak_run)."""
    _need(2)
    rc = launcher.launch([WORKER, "soak"], workers=1, port=23467, timeout=240)
    assert rc == 0


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("mode", PROPS)
def test_ddp_many_gpus_properties(world, mode):
    """BASELINE configs 3/4 at N=2/4/8: rank-ordered gradient sum, bit-identical replicas, decreasing loss, smoke-dist
    send/recv and all_reduce bit-exactness."""
    _need(world)
    rc = launcher.launch([WORKER, mode], workers=world - 1, port=23470 + 4 * world + PROPS.index(mode), timeout=240)
    assert rc == 0


@pytest.mark.parametrize("world", [1, 2, 8])
def test_smoke_dist_payload_main_under_the_launcher(world):
    """examples/smoke-dist/dist_sendrecv.py:41-56 is what the operator's e2e launches (1 Master + N Workers,
    test/e2e/v1/default/defaults.go:121-153) and judges by exit code: run the drop-in's main() the same way
    (world 1 = a Master alone, which also runs on a one-GPU box)."""
    _need(world)
    rc = launcher.launch(["-m", "pytorch_operator_b200.dist_sendrecv"], workers=world - 1, port=23520 + world,
                         timeout=240)
    assert rc == 0


def test_mnist_payload_ddp_under_the_launcher():
    """The mnist drop-in as a 1 Master + 1 Worker job with the nccl manifest's arguments."""
    _need(2)
    rc = launcher.launch(["-m", "pytorch_operator_b200.mnist", "--synthetic", "30", "--test-batch-size", "100",
                          "--backend", "nccl"], workers=1, port=23540, timeout=300)
    assert rc == 0
