"""Single-box launcher that reproduces the operator's env contract.

``cluster_spec`` restates ``setClusterSpec`` (pkg/controller.v1/pytorch/pod.go:234-281; golden
values in pod_test.go:108-150): for a job with 1 Master and ``workers`` Workers it yields, per
replica, MASTER_PORT (default container port 23456, constants.go:31), MASTER_ADDR
(``localhost`` on the master, ``<job>-master-0`` on workers), WORLD_SIZE (= total replicas),
RANK (0 for the master, index+1 for a worker) and PYTHONUNBUFFERED=0.  ``launch`` starts one
process per replica on THIS box with exactly that environment, except that MASTER_ADDR is
resolved to 127.0.0.1 (there is no headless Service DNS outside a cluster) and each process
is pinned to one GPU with CUDA_VISIBLE_DEVICES -- the single-GPU-per-pod view the device
plugin gives a pod (``nvidia.com/gpu: 1`` in examples/mnist/v1/pytorch_job_mnist_nccl.yaml).

    python -m pytorch_operator_b200.launcher --workers 7 -- -m pytorch_operator_b200.mnist --backend nccl --synthetic 100
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys

DEFAULT_PORT = 23456  # pkg/apis/pytorch/v1/constants.go:31


def cluster_spec(job_name: str, workers: int, rtype: str, index: int, port: int = DEFAULT_PORT) -> dict:
    rtype = rtype.lower()
    if rtype not in ("master", "worker"):
        raise ValueError("replica type must be Master or Worker")
    if rtype == "master" and index != 0:
        raise ValueError("invalid config: There should be only a single master with index=0")  # pod.go:247-249
    total = 1 + workers
    if rtype == "master":
        addr, rank = "localhost", 0
    else:
        addr, rank = f"{job_name}-master-0", index + 1
    return {"MASTER_PORT": str(port), "MASTER_ADDR": addr, "WORLD_SIZE": str(total), "RANK": str(rank),
            "PYTHONUNBUFFERED": "0"}


def launch(py_args, workers: int, job_name: str = "local-pytorchjob", port: int = DEFAULT_PORT, gpus=None,
           timeout=None) -> int:
    procs = []
    total = 1 + workers
    for r in range(total):
        env = dict(os.environ)
        env.update(cluster_spec(job_name, workers, "master" if r == 0 else "worker", max(r - 1, 0), port))
        env["MASTER_ADDR"] = "127.0.0.1"
        if gpus is not False:
            env["CUDA_VISIBLE_DEVICES"] = str((gpus or list(range(total)))[r])
        procs.append(subprocess.Popen([sys.executable, *py_args], env=env))
    rc = 0
    for p in procs:
        try:
            rc = p.wait(timeout=timeout) or rc
        except subprocess.TimeoutExpired:
            p.kill()
            rc = rc or 124
    return rc  # the job succeeds iff every replica exits 0 (status.go:89-144)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--workers", type=int, default=1)
    ap.add_argument("--port", type=int, default=DEFAULT_PORT)
    ap.add_argument("--job-name", default="local-pytorchjob")
    ap.add_argument("--timeout", type=float, default=None)
    ap.add_argument("payload", nargs=argparse.REMAINDER, help="-- <python args of the payload>")
    a = ap.parse_args(argv)
    payload = a.payload[1:] if a.payload[:1] == ["--"] else a.payload
    sys.exit(launch(payload, a.workers, a.job_name, a.port, timeout=a.timeout))


if __name__ == "__main__":
    main()
