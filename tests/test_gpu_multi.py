"""Two-GPU parity (BASELINE configs 3 and 4 at N=2): flat-bucket DDP over NCCL against the fixture the
reference produced with DistributedDataParallelCPU on gloo, and the smoke-dist send/recv + all_reduce
bit-exactness check.  Skipped when fewer than two GPUs are visible."""
import os
import sys

import pytest
import torch

from pytorch_operator_b200 import launcher

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("mode", ["eager", "graph"])
def test_ddp_two_gpus_matches_reference_fixture(mode):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = 23456 if mode == "eager" else 23457
    rc = launcher.launch([os.path.join(ROOT, "tests", "ddp_worker.py"), mode], workers=1, port=port, timeout=120)
    assert rc == 0
