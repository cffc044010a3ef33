"""``SGD`` -- optim.SGD(lr, momentum) of the reference (mnist.py:140) as one fused kernel
over the flat parameter bucket, with DDP's 1/world_size folded in as ``grad_scale``."""
from __future__ import annotations

import torch

from . import _cabi
from .net import Net, _stream


class SGD:
    """Same constructor shape as ``torch.optim.SGD(params, lr, momentum)`` for the options the
    reference uses (dampening 0, no nesterov, no weight decay).  ``params`` may be
    ``model.parameters()`` of a ``Net`` (or of our ``DistributedDataParallel`` around it) --
    the optimizer then updates the model's flat bucket in one launch."""

    def __init__(self, params, lr=0.01, momentum=0.0, model: Net | None = None):
        params = list(params)
        if model is None and params:
            ref = getattr(params[0], "_b200_net", None)
            model = ref() if ref is not None else None
        if model is None:
            raise ValueError("the parameters must come from a pytorch_operator_b200.Net (or pass model=)")
        mine = {p.data_ptr() for p in model.parameters()}
        if {p.data_ptr() for p in params} != mine:
            raise ValueError("SGD must own exactly the parameters of the given Net")
        self.model = model
        self.lr, self.momentum = float(lr), float(momentum)
        self.grad_scale = 1.0  # DistributedDataParallel sets 1/world_size
        self.first_step = True
        self.param_groups = [{"params": params, "lr": self.lr, "momentum": self.momentum}]

    def zero_grad(self, set_to_none: bool = False):
        # mnist.py:39.  The backward kernels overwrite the gradient bucket, so nothing to clear.
        return None

    def effective_grad_scale(self) -> float:
        """1/world_size when the model is wrapped in DistributedDataParallel (the bucket then
        holds the SUM over ranks), else the user-set ``grad_scale``."""
        ddp = self.model._ddp
        return 1.0 / ddp.world_size if ddp is not None else self.grad_scale

    @torch.no_grad()
    def step(self):
        m = self.model
        self.grad_scale = self.effective_grad_scale()
        n = m.flat_params.numel()
        L = _cabi.lib()
        _cabi.check(L.b200mnist_sgd_momentum(m.flat_params.data_ptr(), m.flat_grads.data_ptr(),
                                             m.flat_momentum.data_ptr(), n, self.lr, self.momentum, self.grad_scale,
                                             1 if self.first_step else 0, _stream()), "sgd_momentum")
        self.first_step = False
