"""Drop-in for the payload the operator launches: examples/mnist/mnist.py of the reference.

Same CLI (mnist.py:78-102), same symbols (``Net``, ``train``, ``test``, ``should_distribute``,
``is_distributed``, ``main``), same stdout lines (mnist.py:45-47,64), same env contract
(``WORLD_SIZE`` read at import, mnist.py:14; ``env://`` rendezvous from MASTER_ADDR /
MASTER_PORT / RANK / WORLD_SIZE as injected by pkg/controller.v1/pytorch/pod.go:234-281) --
but the training step runs on the sm_100a kernels of libb200mnist.so with one flat-bucket
all_reduce per step.  Additive flags only: ``--synthetic N`` (seeded synthetic batches
instead of the MNIST download, which needs a network), ``--data-root``, ``--no-graph``.

    python -m pytorch_operator_b200.mnist --backend nccl --synthetic 200
"""
from __future__ import print_function

import argparse
import logging
import os
import sys

import torch
import torch.distributed as dist

from .data import DeviceMNIST, find_mnist
from .ddp import DistributedDataParallel
from .net import Net, StepPipeline
from .optim import SGD

try:  # mnist.py:6 imports tensorboardX unconditionally; it is optional here
    from tensorboardX import SummaryWriter
except Exception:  # pragma: no cover - depends on the image
    class SummaryWriter(object):
        """No-op stand-in with the two calls the payload makes (mnist.py:49,65,108)."""

        def __init__(self, *a, **k):
            pass

        def add_scalar(self, *a, **k):
            pass

WORLD_SIZE = int(os.environ.get('WORLD_SIZE', 1))


def train(args, model, device, train_loader, optimizer, epoch, writer):
    """mnist.py:35-49.  Per batch: H2D copy, fused zero_grad/forward/nll_loss/backward/
    all_reduce/step (CUDA-graph replays fed through a device ring once the batch shape is known), and the same
    ``loss.item()`` / print / add_scalar every ``log_interval`` batches."""
    model.train()
    graphs = getattr(model, "_b200_step_graphs", None)
    if graphs is None:
        graphs = model._b200_step_graphs = {}
    use_graph = not getattr(args, "no_graph", False)
    for batch_idx, (data, target) in enumerate(train_loader):
        B = len(data)
        g = graphs.get(B)
        if g is None and use_graph and not optimizer.first_step:
            g = graphs[B] = StepPipeline(model, optimizer, B, unroll=4)
        for other in graphs.values():   # steps run in submission order: a pipeline of another batch size launches first
            if other is not g:
                other.flush()
        if g is not None:
            loss = g.submit(data, target)  # H2D on a copy stream, overlapped with the previous steps
        else:
            data, target = data.to(device, non_blocking=True), target.to(device, non_blocking=True)
            loss = model.train_step(data, target, optimizer)
        if batch_idx % args.log_interval == 0:
            print('Train Epoch: {} [{}/{} ({:.0f}%)]\tloss={:.4f}'.format(
                epoch, batch_idx * len(data), len(train_loader.dataset),
                100. * batch_idx / len(train_loader), loss.item()))
            niter = epoch * len(train_loader) + batch_idx
            writer.add_scalar('loss', loss.item(), niter)
    for g in graphs.values():
        g.flush()


def test(args, model, device, test_loader, writer, epoch):
    """mnist.py:51-65.  The per-batch NLL sum and correct count are reduced on the device;
    one D2H read at the end instead of two ``.item()`` syncs per batch."""
    model.eval()
    totals = torch.zeros(2, dtype=torch.float64, device=device)
    with torch.no_grad():
        for data, target in test_loader:
            data, target = data.to(device, non_blocking=True), target.to(device, non_blocking=True)
            totals += model.evaluate(data, target).double()
    test_loss, correct = totals.tolist()
    test_loss /= len(test_loader.dataset)
    print('\naccuracy={:.4f}\n'.format(float(correct) / len(test_loader.dataset)))
    writer.add_scalar('accuracy', float(correct) / len(test_loader.dataset), epoch)


def should_distribute():
    return dist.is_available() and WORLD_SIZE > 1


def is_distributed():
    return dist.is_available() and dist.is_initialized()


class SyntheticLoader(list):
    """A list of seeded synthetic (data, target) batches in pinned memory, with the
    ``.dataset`` length train()/test() print.  Rank-dependent seeds so the all_reduce is
    not trivial (the reference, having no DistributedSampler, feeds every rank the same data)."""

    def __init__(self, nbatches, batch_size, seed, pin=True):
        g = torch.Generator().manual_seed(seed)
        batches = []
        for _ in range(nbatches):
            x = torch.randn(batch_size, 1, 28, 28, generator=g)
            t = torch.randint(0, 10, (batch_size,), generator=g)
            if pin and torch.cuda.is_available():
                x, t = x.pin_memory(), t.pin_memory()
            batches.append((x, t))
        super().__init__(batches)
        self.dataset = range(nbatches * batch_size)


def pick_device(use_cuda):
    """mnist.py:112 uses bare "cuda": one GPU per pod.  On a single box with several visible
    GPUs (no device plugin) map RANK (or LOCAL_RANK) onto them."""
    if not use_cuda:
        return torch.device("cpu")
    n = torch.cuda.device_count()
    idx = int(os.environ.get("LOCAL_RANK", os.environ.get("RANK", 0))) % max(n, 1)
    torch.cuda.set_device(idx)
    return torch.device("cuda", idx)


def main(argv=None):
    # Training settings
    parser = argparse.ArgumentParser(description='PyTorch MNIST Example')
    parser.add_argument('--batch-size', type=int, default=64, metavar='N',
                        help='input batch size for training (default: 64)')
    parser.add_argument('--test-batch-size', type=int, default=1000, metavar='N',
                        help='input batch size for testing (default: 1000)')
    parser.add_argument('--epochs', type=int, default=1, metavar='N',
                        help='number of epochs to train (default: 10)')
    parser.add_argument('--lr', type=float, default=0.01, metavar='LR',
                        help='learning rate (default: 0.01)')
    parser.add_argument('--momentum', type=float, default=0.5, metavar='M',
                        help='SGD momentum (default: 0.5)')
    parser.add_argument('--no-cuda', action='store_true', default=False,
                        help='disables CUDA training')
    parser.add_argument('--seed', type=int, default=1, metavar='S',
                        help='random seed (default: 1)')
    parser.add_argument('--log-interval', type=int, default=10, metavar='N',
                        help='how many batches to wait before logging training status')
    parser.add_argument('--save-model', action='store_true', default=False,
                        help='For Saving the current Model')
    parser.add_argument('--dir', default='logs', metavar='L',
                        help='directory where summary logs are stored')
    if dist.is_available():
        parser.add_argument('--backend', type=str, help='Distributed backend',
                            choices=[dist.Backend.GLOO, dist.Backend.NCCL, dist.Backend.MPI],
                            default=dist.Backend.GLOO)
    # additive flags
    parser.add_argument('--synthetic', type=int, default=0, metavar='NB',
                        help='train on NB seeded synthetic batches instead of downloading MNIST')
    parser.add_argument('--data-root', default='../data', metavar='D',
                        help="where torchvision keeps MNIST (<D>/MNIST/raw/*-ubyte[.gz]); the reference hard-codes '../data'")
    parser.add_argument('--no-graph', action='store_true', default=False,
                        help='do not capture the training step in a CUDA graph')
    args = parser.parse_args(argv)
    logging.basicConfig(level=logging.INFO, stream=sys.stderr)  # stdout stays the reference's lines only
    use_cuda = not args.no_cuda and torch.cuda.is_available()
    if use_cuda:
        print('Using CUDA')
    else:
        raise SystemExit('this payload runs on the CUDA kernels of libb200mnist.so only; no GPU is visible '
                         '(there is no CPU fallback)')

    writer = SummaryWriter(args.dir)

    torch.manual_seed(args.seed)

    device = pick_device(use_cuda)

    if should_distribute():
        print('Using distributed PyTorch with {} backend'.format(args.backend))
        dist.init_process_group(backend=args.backend)

    rank = dist.get_rank() if is_distributed() else 0
    if args.synthetic > 0:
        train_loader = SyntheticLoader(args.synthetic, args.batch_size, 1234 + rank)
        test_loader = SyntheticLoader(2, args.test_batch_size, 4321)
    elif find_mnist(args.data_root, True) and find_mnist(args.data_root, False):
        # the idx files are already on disk: keep the dataset on the GPU (no DataLoader process, no per-step H2D)
        train_loader = DeviceMNIST.from_files(*find_mnist(args.data_root, True), batch_size=args.batch_size,
                                              device=device, shuffle=True, seed=args.seed)
        test_loader = DeviceMNIST.from_files(*find_mnist(args.data_root, False), batch_size=args.test_batch_size,
                                             device=device, shuffle=False)
    else:
        from torchvision import datasets, transforms
        kwargs = {'num_workers': 1, 'pin_memory': True}
        tf = transforms.Compose([transforms.ToTensor(), transforms.Normalize((0.1307,), (0.3081,))])
        train_loader = torch.utils.data.DataLoader(
            datasets.MNIST('../data', train=True, download=True, transform=tf),
            batch_size=args.batch_size, shuffle=True, **kwargs)
        test_loader = torch.utils.data.DataLoader(
            datasets.MNIST('../data', train=False, transform=tf),
            batch_size=args.test_batch_size, shuffle=False, **kwargs)

    model = Net().to(device)

    if is_distributed():
        model = DistributedDataParallel(model)

    optimizer = SGD(model.parameters(), lr=args.lr, momentum=args.momentum)

    for epoch in range(1, args.epochs + 1):
        train(args, model, device, train_loader, optimizer, epoch, writer)
        test(args, model, device, test_loader, writer, epoch)

    if (args.save_model):
        if rank == 0:  # the reference lets every rank race on this path (mnist.py:146-147)
            torch.save({k: v.detach().cpu().clone() for k, v in model.state_dict().items()}, "mnist_cnn.pt")

    if is_distributed():
        model.release_graphs()  # graphs that captured NCCL kernels must go before the process group
        getattr(model, "_b200_step_graphs", {}).clear()
        model.close()           # unmap / free the peer windows of the fused exchange (no-op on the NCCL path)
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
