"""B200-native drop-in for the data-parallel payloads jiaqianjing/pytorch-operator launches
(examples/mnist and examples/smoke-dist).  The compute runs in libb200mnist.so (hand-written
sm_100a CUDA kernels behind the C ABI of include/b200mnist.h); this package is the thin
Python host side that mirrors the reference payloads' interface."""
from . import _cabi
from .ddp import DistributedDataParallel
from .net import NAMES, InputRing, Net, StepGraph, StepPipeline
from .optim import SGD

__all__ = ["Net", "SGD", "DistributedDataParallel", "StepGraph", "StepPipeline", "InputRing", "NAMES", "_cabi"]
