"""Two-GPU parity (BASELINE configs 3 and 4 at N=2): flat-bucket DDP -- NCCL all_reduce path and the fused peer-memory all-reduce+SGD kernel -- against the fixture the
reference produced with DistributedDataParallelCPU on gloo, and the smoke-dist send/recv + all_reduce
bit-exactness check.  Skipped when fewer than two GPUs are visible."""
import os
import sys

import pytest
import torch

from pytorch_operator_b200 import launcher

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("mode", ["eager", "graph", "fused", "fused-graph"])
def test_ddp_two_gpus_matches_reference_fixture(mode):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = 23456 + ["eager", "graph", "fused", "fused-graph"].index(mode)
    rc = launcher.launch([os.path.join(ROOT, "tests", "ddp_worker.py"), mode], workers=1, port=port, timeout=120,
                         gpus=False if mode.startswith("fused") else None)
    assert rc == 0


@pytest.mark.parametrize("world", [4, 8])
@pytest.mark.parametrize("mode", ["prop-nccl", "prop-fused"])
def test_ddp_many_gpus_properties(world, mode):
    """BASELINE configs 3/4 at N=4 and N=8 (no reference fixture beyond ws=2): rank-ordered gradient sum, bit-identical
    replicas, decreasing loss, smoke-dist send/recv and all_reduce bit-exactness."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    rc = launcher.launch([os.path.join(ROOT, "tests", "ddp_worker.py"), mode], workers=world - 1,
                         port=23470 + world + (1 if mode == "prop-fused" else 0), timeout=180,
                         gpus=False if mode == "prop-fused" else None)
    assert rc == 0
