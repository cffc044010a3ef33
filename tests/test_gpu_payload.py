"""The drop-in payloads end to end on one GPU: `python -m pytorch_operator_b200.mnist` with the reference's CLI
and stdout contract (mnist.py:45-47,64), StepPipeline against eager steps, and the env-contract launcher."""
import os
import re
import subprocess
import sys

import pytest
import torch

from oracle import mnist_oracle as mo

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEV = "cuda:0"


def test_payload_cli_runs_and_prints_reference_lines(tmp_path):
    env = dict(os.environ, PYTHONPATH=ROOT, WORLD_SIZE="1")
    env.pop("RANK", None)
    r = subprocess.run([sys.executable, "-m", "pytorch_operator_b200.mnist", "--synthetic", "25", "--batch-size", "64",
                        "--test-batch-size", "500", "--epochs", "1", "--save-model", "--dir", str(tmp_path / "logs")],
                       env=env, cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    out = r.stdout
    assert out.startswith("Using CUDA")
    lines = [ln for ln in out.splitlines() if ln.startswith("Train Epoch")]
    assert len(lines) == 3  # batches 0, 10, 20 of 25 with the default --log-interval 10
    assert re.fullmatch(r"Train Epoch: 1 \[0/1600 \(0%\)\]\tloss=\d\.\d{4}", lines[0]), lines[0]
    assert re.fullmatch(r"Train Epoch: 1 \[640/1600 \(40%\)\]\tloss=\d\.\d{4}", lines[1]), lines[1]
    assert re.search(r"\naccuracy=0\.\d{4}\n", out)
    sd = torch.load(tmp_path / "mnist_cnn.pt")
    assert list(sd) == ["conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias", "fc1.weight", "fc1.bias",
                        "fc2.weight", "fc2.bias"] and sd["fc1.weight"].shape == (500, 800)
    # the first printed loss is the reference's step-0 loss for this seed/batch stream
    import pytorch_operator_b200.mnist as payload
    x, t = payload.SyntheticLoader(1, 64, 1234, pin=False)[0]
    lo, _, _ = mo.loss_and_grads(mo.init_params(1), x, t)
    assert abs(float(lines[0].split("loss=")[1]) - float(lo)) < 1e-4


def test_payload_logged_losses_are_the_current_step(tmp_path):
    """mnist.py:47 prints loss.item() of the step just run.  The pipelined (CUDA-graph) loop writes the loss
    asynchronously into pinned host memory, so every 'loss=' line after batch 0 must wait for its own step:
    the lines of a pipelined run equal those of an eager (--no-graph) run, batch for batch."""
    env = dict(os.environ, PYTHONPATH=ROOT, WORLD_SIZE="1")
    env.pop("RANK", None)
    outs = []
    for extra in ([], ["--no-graph"]):
        r = subprocess.run([sys.executable, "-m", "pytorch_operator_b200.mnist", "--synthetic", "45", "--batch-size", "64",
                            "--test-batch-size", "100", "--log-interval", "4", "--dir", str(tmp_path / "logs"), *extra],
                           env=env, cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append([ln for ln in r.stdout.splitlines() if ln.startswith("Train Epoch")])
    assert len(outs[0]) == 12 and outs[0] == outs[1], (outs[0], outs[1])
    # and they are a real trajectory, not a repeated or zero value
    vals = [float(ln.split("loss=")[1]) for ln in outs[0]]
    assert len(set(vals)) > 6 and min(vals) > 0.0


def test_step_pipeline_matches_eager_steps():
    import pytorch_operator_b200 as pkg
    batches = [mo.synthetic_batch(64, 300 + i) for i in range(7)]

    def run(pipelined):
        torch.manual_seed(1)
        net = pkg.Net().to(DEV)
        opt = pkg.SGD(net.parameters(), lr=0.01, momentum=0.5)
        losses = [float(net.train_step(batches[0][0].to(DEV), batches[0][1].to(DEV), opt))]
        pipe = pkg.StepPipeline(net, opt, 64) if pipelined else None
        for x, t in batches[1:]:
            if pipelined:
                loss = pipe.submit(x.pin_memory(), t.pin_memory())
                pipe.drain()
                assert float(loss) == float(pipe.loss_host[(pipe.steps - 1) & 1])  # the kernel stored it in pinned host memory
            else:
                loss = net.train_step(x.to(DEV), t.to(DEV), opt)
            losses.append(float(loss))
        return losses, net.flat_params.clone()

    l_e, p_e = run(False)
    l_p, p_p = run(True)
    assert l_e == l_p and torch.equal(p_e, p_p)  # same kernels, same order: bit-identical


def test_launcher_runs_single_replica_job():
    from pytorch_operator_b200 import launcher
    rc = launcher.launch(["-m", "pytorch_operator_b200.mnist", "--synthetic", "3", "--test-batch-size", "100",
                          "--backend", "nccl"], workers=0, port=23461, gpus=[0], timeout=300)
    assert rc == 0
