"""Drop-in for examples/smoke-dist/dist_sendrecv.py: the Master <-> Workers send/recv check
the operator's e2e tests launch (test/e2e/v1/default/defaults.go), on the GPU.

Same symbols (``run``, ``init_processes``, ``main``) and the same five log lines
(dist_sendrecv.py:42-54).  Rank 0 sends a 2x2 fp32 tensor to every worker over the process
group (NCCL send/recv over NVLink on the GPU box), the worker squares it with the
``b200mnist_square`` kernel (one IEEE multiply per element, so the result is bit-exact
against ``torch.mul`` on the CPU) and sends it back.  ``run`` additionally VERIFIES what the
reference only logs, and ``allreduce_check`` adds the all_reduce bit-exactness check and
message-size sweep that BASELINE.json asks for (the reference file itself never calls
all_reduce).
"""
import logging
import os
import time

import torch
import torch.distributed as dist

from . import _cabi


def _device():
    if not torch.cuda.is_available():
        raise SystemExit("dist_sendrecv runs on the GPU only (no CPU fallback); no GPU is visible")
    idx = int(os.environ.get("LOCAL_RANK", os.environ.get("RANK", 0))) % torch.cuda.device_count()
    torch.cuda.set_device(idx)
    return torch.device("cuda", idx)


def square(inp):
    """torch.mul(inp, inp) of dist_sendrecv.py:32 on the b200mnist_square kernel."""
    inp = inp.contiguous()
    out = torch.empty_like(inp)
    _cabi.check(_cabi.lib().b200mnist_square(inp.data_ptr(), out.data_ptr(), inp.numel(),
                                             torch.cuda.current_stream().cuda_stream), "square")
    return out


def run():
    """ Simple Send/Recv for testing Master <--> Workers communication """
    rank = dist.get_rank()
    size = dist.get_world_size()
    dev = _device()
    inp = torch.randn(2, 2).to(dev)  # drawn on the CPU generator like dist_sendrecv.py:19
    result = torch.zeros(2, 2, device=dev)
    results = []
    if rank == 0:
        # Send the input tensor to all workers
        for i in range(1, size):
            dist.send(tensor=inp, dst=i)
            # Receive the result tensor from all workers
            dist.recv(tensor=result, src=i)
            logging.info("Result from worker %d : %s", i, result)
            expect = torch.mul(inp.cpu(), inp.cpu())
            if not torch.equal(result.cpu(), expect):
                raise RuntimeError("worker %d returned %s, expected %s" % (i, result.cpu(), expect))
            results.append(result.cpu().clone())
    else:
        # Receive input tensor from master
        dist.recv(tensor=inp, src=0)
        # Elementwise tensor multiplication
        result = square(inp)
        # Send the result tensor back to master
        dist.send(tensor=result, dst=0)
    return inp.cpu(), results


def allreduce_check(sizes=(4, 256, 431296, 1 << 20, 1 << 24), iters=5):
    """all_reduce(SUM) bit-exactness (integer-valued fp32 inputs, so every summation order
    gives the same bits) and bus bandwidth per message size; rank 0 returns the rows."""
    rank, ws, dev = dist.get_rank(), dist.get_world_size(), _device()
    rows = []
    for n in sizes:
        base = (torch.arange(n, dtype=torch.float32) % 251.0)
        buf = (base + rank).to(dev)
        dist.all_reduce(buf)
        expect = base * ws + ws * (ws - 1) / 2
        exact = bool(torch.equal(buf.cpu(), expect))
        buf = torch.ones(n, device=dev)
        for _ in range(2):
            dist.all_reduce(buf)
        torch.cuda.synchronize()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(iters):
            dist.all_reduce(buf)
        t1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([t0.elapsed_time(t1) / iters], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        sec = float(ms) * 1e-3
        rows.append({"bytes": n * 4, "bit_exact": exact, "time_us": sec * 1e6,
                     "busbw_GBps": (n * 4 / sec) * 2 * (ws - 1) / ws / 1e9})
        if not exact:
            raise RuntimeError("all_reduce of %d floats is not bit-exact" % n)
    if rank == 0:
        for r in rows:
            logging.info("all_reduce %(bytes)d B: bit_exact=%(bit_exact)s %(time_us).1f us busBW %(busbw_GBps).2f GB/s", r)
    return rows


def init_processes(fn, backend='gloo'):
    """ Initialize the distributed environment. """
    dist.init_process_group(backend)
    fn()


def main():
    logging.info("Torch version: %s", torch.__version__)

    port = os.environ.get("MASTER_PORT", "{}")
    logging.info("MASTER_PORT: %s", port)

    addr = os.environ.get("MASTER_ADDR", "{}")
    logging.info("MASTER_ADDR: %s", addr)

    world_size = os.environ.get("WORLD_SIZE", "{}")
    logging.info("WORLD_SIZE: %s", world_size)

    rank = os.environ.get("RANK", "{}")
    logging.info("RANK: %s", rank)

    def both():
        run()
        if os.environ.get("B200_SMOKE_ALLREDUCE", "1") == "1":
            allreduce_check()

    # the reference hard-codes gloo (dist_sendrecv.py:36); the GPU box exchanges over NCCL
    init_processes(both, backend=os.environ.get("B200_SMOKE_BACKEND", "nccl"))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    logging.getLogger().setLevel(logging.INFO)
    main()
