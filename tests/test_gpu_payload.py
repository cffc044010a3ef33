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


@pytest.mark.parametrize("unroll", [1, 4])
def test_step_pipeline_matches_eager_steps(unroll):
    """unroll=4: the steps go out four per graph launch; the two left over at the end and a loss somebody reads early
    go out through the single-step graph.  Either way: the eager trajectory, bit for bit."""
    import pytorch_operator_b200 as pkg
    batches = [mo.synthetic_batch(64, 300 + i) for i in range(23)]   # more steps than ring slots: slots get reused

    def run(pipelined):
        torch.manual_seed(1)
        net = pkg.Net().to(DEV)
        opt = pkg.SGD(net.parameters(), lr=0.01, momentum=0.5)
        losses = [float(net.train_step(batches[0][0].to(DEV), batches[0][1].to(DEV), opt))]
        pipe = pkg.StepPipeline(net, opt, 64, unroll=unroll) if pipelined else None
        unread = []
        for k, (x, t) in enumerate(batches[1:]):
            if pipelined:
                unread.append(pipe.submit(x.pin_memory(), t.pin_memory()))
                if unroll == 1:
                    pipe.drain()
                    assert float(unread[-1]) == float(pipe.loss_host[(pipe.steps - 1) % pipe.slots])  # stored in pinned host memory by the kernel
                if k == 9:                      # mid-group read: steps 9 and 10 of the pipeline leave through the single-step graph
                    float(unread[-1])
                if len(unread) == 4:            # read before `slots` more steps are submitted
                    losses += [float(v) for v in unread]
                    unread = []
            else:
                losses.append(float(net.train_step(x.to(DEV), t.to(DEV), opt)))
        if pipelined:
            pipe.drain()
            losses += [float(v) for v in unread]
            assert pipe.steps == 22 and int(pipe.ring.state[1]) == 22
        return losses, net.flat_params.clone()

    l_e, p_e = run(False)
    l_p, p_p = run(True)
    assert l_e == l_p and torch.equal(p_e, p_p)  # same kernels, same order: bit-identical


@pytest.mark.parametrize("packed", [True, False])
def test_resident_input_ring_steps_match_plain_steps(packed):
    """b200mnist_train_step_ring: batches resident in a device ring, read through a device-side cursor (eager and as
    CUDA-graph replays with no copy in between) give bit-identical training to the same batches passed one by one;
    the cursor wraps, and x / target receive the slot's batch."""
    import pytorch_operator_b200 as pkg
    B, slots, steps = 64, 3, 8
    batches = [mo.synthetic_batch(B, 700 + i) for i in range(slots)]

    def plain():
        torch.manual_seed(1)
        net = pkg.Net().to(DEV)
        opt = pkg.SGD(net.parameters(), lr=0.01, momentum=0.5)
        losses = [float(net.train_step(batches[i % slots][0].to(DEV), batches[i % slots][1].to(DEV), opt)) for i in range(steps)]
        return losses, net.flat_params.clone()

    def ringed():
        torch.manual_seed(1)
        net = pkg.Net().to(DEV)
        opt = pkg.SGD(net.parameters(), lr=0.01, momentum=0.5)
        if packed:
            rec = torch.empty(slots, B * 784 * 4 + B * 8, dtype=torch.uint8, device=DEV)
            for i, (x, t) in enumerate(batches):
                rec[i, :B * 784 * 4].view(torch.float32).copy_(x.reshape(-1))
                rec[i, B * 784 * 4:].view(torch.int64).copy_(t)
            ring = pkg.InputRing(B, records=rec)
        else:
            ring = pkg.InputRing(B, x=torch.stack([b[0] for b in batches]).to(DEV), target=torch.stack([b[1] for b in batches]).to(DEV))
        x = torch.zeros(B, 1, 28, 28, device=DEV)
        t = torch.zeros(B, dtype=torch.int64, device=DEV)
        losses = [float(net.train_step(x, t, opt, ring=ring)) for _ in range(2)]        # eager: slots 0, 1
        assert int(ring.cursor) == 2 and torch.equal(x.cpu(), batches[1][0]) and torch.equal(t.cpu(), batches[1][1])
        g = pkg.StepGraph(net, opt, B, x=x, target=t, ring=ring)                         # warm-up leaves the cursor alone
        assert int(ring.cursor) == 2
        for i in range(2, steps):
            losses.append(float(g.replay()))
            assert int(ring.cursor) == (i + 1) % slots
        assert torch.equal(x.cpu(), batches[(steps - 1) % slots][0])
        return losses, net.flat_params.clone()

    (l_p, p_p), (l_r, p_r) = plain(), ringed()
    assert l_p == l_r and torch.equal(p_p, p_r)
    ring = pkg.InputRing(B, x=torch.zeros(2, B, 1, 28, 28, device=DEV), target=torch.zeros(2, B, dtype=torch.int64, device=DEV))
    with pytest.raises(ValueError):
        pkg.Net().to(DEV).train_step(torch.zeros(8, 1, 28, 28, device=DEV), torch.zeros(8, dtype=torch.int64, device=DEV),
                                     None, ring=ring)


def test_launcher_runs_single_replica_job():
    from pytorch_operator_b200 import launcher
    rc = launcher.launch(["-m", "pytorch_operator_b200.mnist", "--synthetic", "3", "--test-batch-size", "100",
                          "--backend", "nccl"], workers=0, port=23461, gpus=[0], timeout=300)
    assert rc == 0
