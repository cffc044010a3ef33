"""``Net`` -- the reference's model (examples/mnist/mnist.py:17-33) on the B200 kernels.

Same class name, same sub-module / parameter names (``conv1.weight`` ... ``fc2.bias``) and
the same initial values for a given ``torch.manual_seed`` as the reference's ``Net``; the
parameters, their gradients and the momentum buffers are views into three flat fp32
buckets on the GPU so that the DDP exchange is one collective and the SGD update one kernel.

Two ways to run a training step, both through libb200mnist.so (no torch math, no CPU
fallback):
  * the reference's own loop shape -- ``output = model(data); F.nll_loss(output, target)
    .backward(); optimizer.step()`` -- via an autograd bridge over the per-layer entry points;
  * ``model.train_step(data, target, optimizer)`` -- the fused step (one C call; captured in
    a CUDA graph by ``StepGraph``), which ``mnist.train()`` uses.
"""
from __future__ import annotations

import weakref
from collections import OrderedDict

import torch
import torch.nn as nn

from . import _cabi

NAMES = ["conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias",
         "fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias"]
SHAPES = [(20, 1, 5, 5), (20,), (50, 20, 5, 5), (50,), (500, 800), (500,), (10, 500), (10,)]


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def flat_layout():
    """[(name, shape, offset, numel)] of the flat bucket, as the library lays it out."""
    L = _cabi.lib()
    return [(n, s, int(L.b200mnist_tensor_offset(i)), int(L.b200mnist_tensor_elems(i)))
            for i, (n, s) in enumerate(zip(NAMES, SHAPES))]


class InputRing:
    """Batches that already sit in device memory plus a device-side cursor (``b200mnist_ring_t``): a step called with
    ``ring=`` reads the slot under the cursor in its first kernel and advances the cursor in its last one, so a loop
    of CUDA-graph replays needs no per-step copy and no host-side pointer (what the reference does with a DataLoader
    and ``data.to(device)``, mnist.py:37-38, when the data fits in HBM).

    ``records``: uint8 [slots, row] with row >= B*784*4 + B*8 and a multiple of 16: each row one packed batch (fp32
    images, then int64 targets) -- or pass ``x`` [slots,B,1,28,28] float32 and ``target`` [slots,B] int64.
    ``host_fed``: the slots are refilled by a copy stream while the steps run (``StepPipeline``): adds the per-slot
    ``ready`` sequence numbers the steps wait on, the ``steps_done`` counter and its pinned-host mirror."""

    def __init__(self, batch, records=None, x=None, target=None, host_fed=False):
        import ctypes
        self.batch = batch
        if records is not None:
            xb = batch * 784 * 4
            if records.dtype != torch.uint8 or records.dim() != 2 or records.shape[1] < xb + batch * 8 \
                    or not records.is_contiguous() or records.shape[1] % 16:
                raise ValueError("records must be contiguous uint8 [slots, >= B*784*4 + B*8] with rows a multiple of 16 bytes")
            self._keep = (records,)
            x_ptr, t_ptr = records.data_ptr(), records.data_ptr() + xb
            x_stride, t_stride, slots, dev = records.shape[1] // 4, records.shape[1] // 8, records.shape[0], records.device
        else:
            if x.dtype != torch.float32 or target.dtype != torch.int64 or not x.is_contiguous() \
                    or not target.is_contiguous() or tuple(x.shape[1:]) != (batch, 1, 28, 28) \
                    or tuple(target.shape) != (x.shape[0], batch):
                raise ValueError("x must be float32 [slots,B,1,28,28] and target int64 [slots,B], both contiguous")
            self._keep = (x, target)
            x_ptr, t_ptr, x_stride, t_stride, slots, dev = x.data_ptr(), target.data_ptr(), batch * 784, batch, x.shape[0], x.device
        self.slots, self.device, self.host_fed = slots, dev, host_fed
        self.state = torch.zeros(2, dtype=torch.int32, device=dev)     # [cursor, steps done]
        self.cursor = self.state[0:1]
        self.landed = torch.zeros(1, dtype=torch.int32, device=dev) if host_fed else None
        self.report_host = torch.zeros(slots, dtype=torch.int64).pin_memory() if host_fed else None
        self.desc = _cabi.Ring(x_ptr, t_ptr, x_stride, t_stride, slots, self.state.data_ptr(),
                               self.landed.data_ptr() if host_fed else None,
                               self.report_host.data_ptr() if host_fed else None)
        self.desc_ptr = ctypes.addressof(self.desc)
        self._report = (ctypes.c_uint64 * slots).from_address(self.report_host.data_ptr()) if host_fed else None

    def seek(self, slot: int):
        """Point the cursor at ``slot`` (stream-ordered)."""
        self.cursor.fill_(int(slot) % self.slots)

    def report(self, step: int):
        """Host-fed rings: (done, loss) of 0-based step ``step`` as last reported to pinned host memory by the step's
        second-to-last kernel (one 8-byte word: no CUDA call, no tearing).  ``done`` turns True when the step has reported;
        ``loss`` is None once the slot has been reused by a later step."""
        word = self._report[step % self.slots]
        seq = word >> 32
        if seq < step + 1:
            return False, None
        if seq > step + 1:
            return True, None
        import struct
        return True, struct.unpack("<f", struct.pack("<I", word & 0xFFFFFFFF))[0]

    # a StepGraph runs one real step on the ring while warming up; these put the ring back exactly as it was
    def _snapshot(self):
        torch.cuda.synchronize(self.device)
        snap = {"state": self.state.clone()}
        if self.host_fed:
            snap.update(landed=self.landed.clone(), report=self.report_host.clone())
            self.landed.fill_(int(self.state[1]) + 1)    # let the warm-up step find "its" batch
        return snap

    def _restore(self, snap):
        torch.cuda.synchronize(self.device)
        self.state.copy_(snap["state"])
        if self.host_fed:
            self.landed.copy_(snap["landed"])
            self.report_host.copy_(snap["report"])
        torch.cuda.synchronize(self.device)


class _Leaf(nn.Module):
    """Holds ``weight`` and ``bias`` so state_dict keys match the reference's layers."""

    def __init__(self):
        super().__init__()
        self.register_parameter("weight", None)
        self.register_parameter("bias", None)


class Net(nn.Module):
    def __init__(self):
        super().__init__()
        # Build the reference's layers once on the CPU, in its declaration order
        # (mnist.py:20-23), so the default init consumes the RNG stream identically.
        ref_layers = [nn.Conv2d(1, 20, 5, 1), nn.Conv2d(20, 50, 5, 1), nn.Linear(4 * 4 * 50, 500), nn.Linear(500, 10)]
        n = int(_cabi.lib().b200mnist_flat_elems())
        flat = torch.zeros(n, dtype=torch.float32)
        lay = flat_layout()
        with torch.no_grad():
            for li, layer in enumerate(ref_layers):
                for pi, p in enumerate((layer.weight, layer.bias)):
                    _, _, off, cnt = lay[2 * li + pi]
                    flat[off:off + cnt] = p.reshape(-1)
        self.conv1, self.conv2, self.fc1, self.fc2 = _Leaf(), _Leaf(), _Leaf(), _Leaf()
        self._layout = lay
        self._ddp = None          # set by ddp.DistributedDataParallel
        self._ws = {}             # batch size -> workspace tensor
        self._loss = None
        self._graphs = []         # weak refs to the StepGraphs captured on this model
        self._bind(flat, torch.zeros_like(flat), torch.zeros_like(flat))

    # ---- flat buckets and their named views -------------------------------------
    def _bind(self, flat_p, flat_g, flat_m):
        self.flat_params, self.flat_grads, self.flat_momentum = flat_p, flat_g, flat_m
        leaves = (self.conv1, self.conv1, self.conv2, self.conv2, self.fc1, self.fc1, self.fc2, self.fc2)
        for (name, shape, off, cnt), leaf in zip(self._layout, leaves):
            attr = name.split(".")[1]
            p = nn.Parameter(flat_p[off:off + cnt].view(shape), requires_grad=True)
            p.grad = flat_g[off:off + cnt].view(shape)
            p._b200_net = weakref.ref(self)  # lets optim.SGD(model.parameters(), ...) find the bucket
            leaf._parameters[attr] = p
        self._ws = {}
        self._loss = torch.zeros(2, dtype=torch.float32, device=flat_p.device)

    def _apply(self, fn, recurse=True):
        # .to(device)/.cuda() must move the buckets and re-create the views, not the views.
        if self._ddp is not None and getattr(self._ddp, "fused", False):
            raise RuntimeError("this Net's buckets live in a peer window of the fused DDP exchange: call "
                               "DistributedDataParallel.close() before moving or casting the model")
        self._bind(fn(self.flat_params), fn(self.flat_grads), fn(self.flat_momentum))
        return self

    @property
    def device(self):
        return self.flat_params.device

    def release_graphs(self):
        """Destroy every CUDA graph captured on this model (call before destroy_process_group())."""
        for ref in self._graphs:
            g = ref()
            if g is not None:
                g.release()
        self._graphs = []

    def named_views(self, flat):
        return OrderedDict((name, flat[off:off + cnt].view(shape)) for name, shape, off, cnt in self._layout)

    def load_named(self, named):
        """Copy {name: tensor} (e.g. a reference state_dict or oracle params) into the bucket."""
        with torch.no_grad():
            for name, v in self.named_views(self.flat_params).items():
                src = named[name] if name in named else named["module." + name]
                v.copy_(torch.as_tensor(src).to(v.device, torch.float32).reshape(v.shape))

    def workspace(self, batch: int) -> torch.Tensor:
        ws = self._ws.get(batch)
        if ws is None:
            nbytes = int(_cabi.lib().b200mnist_workspace_bytes(batch))
            if nbytes < 0:
                raise ValueError(f"unsupported batch size {batch}")
            ws = torch.zeros(nbytes, dtype=torch.uint8, device=self.device)  # zero-filled once (ticket counters)
            self._ws[batch] = ws
        return ws

    def _check_inputs(self, x, target=None):
        if self.device.type != "cuda":
            raise RuntimeError("Net runs on the CUDA kernels of libb200mnist.so only: move it with .to('cuda') "
                               "(there is no CPU fallback)")
        if x.device != self.device or x.dtype != torch.float32 or tuple(x.shape[1:]) != (1, 28, 28):
            raise ValueError("data must be float32 [B,1,28,28] on the model's device")
        if target is not None and (target.device != self.device or target.dtype != torch.int64
                                   or target.shape != (x.shape[0],)):
            raise ValueError("target must be int64 [B] on the model's device")
        return x.contiguous()

    # ---- reference-shaped forward (mnist.py:25-33) ---------------------------------
    def forward(self, x):
        x = self._check_inputs(x)
        if torch.is_grad_enabled() and self.training:
            params = [getattr(getattr(self, n.split(".")[0]), n.split(".")[1]) for n in NAMES]
            return _LeNetFunction.apply(self, x, *params)
        B = x.shape[0]
        logp = torch.empty(B, 10, dtype=torch.float32, device=self.device)
        dummy_t = torch.zeros(B, dtype=torch.int64, device=self.device)
        ws = self.workspace(B)
        L = _cabi.lib()
        _cabi.check(L.b200mnist_forward_eval(self.flat_params.data_ptr(), x.data_ptr(), dummy_t.data_ptr(), B,
                                             ws.data_ptr(), ws.numel(), logp.data_ptr(), self._loss.data_ptr(),
                                             _stream()), "forward_eval")
        return logp

    # ---- fused entry points -----------------------------------------------------
    def forward_backward(self, x, target, logp_out=None, loss_out=None, grads_out=None):
        """mnist.py:39-42 in one call: fills ``flat_grads`` (this rank's gradient) and returns
        the mean-NLL loss as a 0-d device tensor (no sync).  ``loss_out``: optional fp32 device
        tensor (>= 1 element) to receive the loss instead of the model's own scalar."""
        x = self._check_inputs(x, target)
        B = x.shape[0]
        ws = self.workspace(B)
        L = _cabi.lib()
        lo = self._loss if loss_out is None else loss_out
        gr = self.flat_grads if grads_out is None else grads_out  # e.g. the staging bucket of DDP's peer window
        _cabi.check(L.b200mnist_forward_backward(self.flat_params.data_ptr(), gr.data_ptr(),
                                                 x.data_ptr(), target.data_ptr(), B, ws.data_ptr(), ws.numel(),
                                                 lo.data_ptr(), _cabi.ptr(logp_out), _stream()),
                    "forward_backward")
        return lo[0]

    def train_step(self, x, target, optimizer, loss_out=None, ring=None):
        """mnist.py:39-43 in one call.  With DDP attached the gradient bucket is all-reduced
        between the backward and the update, and 1/world_size is folded into the update.
        ``ring`` (an ``InputRing``): the batch comes from the ring slot under its cursor; ``x`` / ``target`` are then
        the buffers the step leaves that batch in."""
        if ring is not None and (ring.batch != x.shape[0] or ring.device != self.device):
            raise ValueError("the ring holds batches of another size or on another device")
        if self._ddp is None:
            lo = self._loss if loss_out is None else loss_out
            x = self._check_inputs(x, target)
            B = x.shape[0]
            ws = self.workspace(B)
            L = _cabi.lib()
            tail = (B, ws.data_ptr(), ws.numel(), optimizer.lr, optimizer.momentum, 1 if optimizer.first_step else 0,
                    lo.data_ptr(), _stream())
            head = (self.flat_params.data_ptr(), self.flat_grads.data_ptr(), self.flat_momentum.data_ptr())
            if ring is None:
                _cabi.check(L.b200mnist_train_step(*head, x.data_ptr(), target.data_ptr(), *tail), "train_step")
            else:
                _cabi.check(L.b200mnist_train_step_ring(*head, ring.desc_ptr, x.data_ptr(), target.data_ptr(), *tail),
                            "train_step_ring")
            optimizer.first_step = False
            return lo[0]
        if self._ddp.fused:
            x = self._check_inputs(x, target)
            return self._ddp.fused_train_step(x, target, optimizer, loss_out, ring)
        if ring is not None:
            raise RuntimeError("a resident input ring needs the single-process or the fused DDP step")
        loss = self.forward_backward(x, target, loss_out=loss_out)
        self._ddp.reduce_gradients()
        optimizer.step()
        return loss

    def evaluate(self, x, target, stats_out=None, logp_out=None):
        """One batch of test() (mnist.py:56-61): returns a device tensor [nll_sum, n_correct]."""
        x = self._check_inputs(x, target)
        B = x.shape[0]
        ws = self.workspace(B)
        if stats_out is None:
            stats_out = torch.zeros(2, dtype=torch.float32, device=self.device)
        L = _cabi.lib()
        _cabi.check(L.b200mnist_forward_eval(self.flat_params.data_ptr(), x.data_ptr(), target.data_ptr(), B,
                                             ws.data_ptr(), ws.numel(), _cabi.ptr(logp_out), stats_out.data_ptr(),
                                             _stream()), "forward_eval")
        return stats_out


class _LeNetFunction(torch.autograd.Function):
    """Autograd bridge: the whole Net as one node, forward and backward on the per-layer
    entry points.  Parameter gradients are written straight into the flat gradient bucket
    (overwritten, as after the reference's zero_grad(); not accumulated)."""

    @staticmethod
    def forward(ctx, net, x, *params):
        L, s, dev, B = _cabi.lib(), _stream(), x.device, x.shape[0]
        P = net.named_views(net.flat_params)
        f32 = dict(dtype=torch.float32, device=dev)
        p1 = torch.empty(B, 20, 12, 12, **f32)
        i1 = torch.empty(B, 20, 12, 12, dtype=torch.uint8, device=dev)
        p2 = torch.empty(B, 50, 4, 4, **f32)
        i2 = torch.empty(B, 50, 4, 4, dtype=torch.uint8, device=dev)
        h = torch.empty(B, 500, **f32)
        logp = torch.empty(B, 10, **f32)
        scratch = torch.zeros(int(L.b200mnist_fc1_scratch_bytes(B)) // 4, **f32)
        _cabi.check(L.b200mnist_conv1_relu_pool_fwd(x.data_ptr(), P["conv1.weight"].data_ptr(),
                                                    P["conv1.bias"].data_ptr(), p1.data_ptr(), i1.data_ptr(), B, s),
                    "conv1 fwd")
        _cabi.check(L.b200mnist_conv2_relu_pool_fwd(p1.data_ptr(), P["conv2.weight"].data_ptr(),
                                                    P["conv2.bias"].data_ptr(), p2.data_ptr(), i2.data_ptr(), B, s),
                    "conv2 fwd")
        _cabi.check(L.b200mnist_fc1_relu_fwd(p2.data_ptr(), P["fc1.weight"].data_ptr(), P["fc1.bias"].data_ptr(),
                                             h.data_ptr(), B, scratch.data_ptr(), s), "fc1 fwd")
        # fc2 + log_softmax through the head kernel (targets unused here: zeros)
        t0 = torch.zeros(B, dtype=torch.int64, device=dev)
        hs = torch.zeros(B + 64, **f32)
        loss = torch.empty(1, **f32)
        _cabi.check(L.b200mnist_fc2_logsoftmax_nll(h.data_ptr(), P["fc2.weight"].data_ptr(), P["fc2.bias"].data_ptr(),
                                                   t0.data_ptr(), B, logp.data_ptr(), loss.data_ptr(), 0, 0,
                                                   hs.data_ptr(), s), "head fwd")
        ctx.net = net
        ctx.save_for_backward(x, p1, i1, p2, i2, h, logp)
        return logp

    @staticmethod
    def backward(ctx, dlogp):
        net = ctx.net
        x, p1, i1, p2, i2, h, logp = ctx.saved_tensors
        L, s, dev, B = _cabi.lib(), _stream(), x.device, x.shape[0]
        P, G = net.named_views(net.flat_params), net.named_views(net.flat_grads)
        f32 = dict(dtype=torch.float32, device=dev)
        dlogp = dlogp.contiguous().to(torch.float32)
        dz2, dz1 = torch.empty(B, 10, **f32), torch.empty(B, 500, **f32)
        dp2, dp1 = torch.empty(B, 800, **f32), torch.empty(B, 20, 12, 12, **f32)
        _cabi.check(L.b200mnist_head_bwd(logp.data_ptr(), dlogp.data_ptr(), h.data_ptr(), P["fc2.weight"].data_ptr(),
                                         B, dz2.data_ptr(), dz1.data_ptr(), s), "head bwd")
        sc = torch.empty(int(L.b200mnist_fc_bwd_scratch_bytes(B)), dtype=torch.uint8, device=dev)
        _cabi.check(L.b200mnist_fc_bwd(dz2.data_ptr(), dz1.data_ptr(), h.data_ptr(), p2.data_ptr(),
                                       P["fc1.weight"].data_ptr(), B, G["fc2.weight"].data_ptr(),
                                       G["fc2.bias"].data_ptr(), G["fc1.weight"].data_ptr(), G["fc1.bias"].data_ptr(),
                                       dp2.data_ptr(), sc.data_ptr(), s), "fc bwd")
        sc2 = torch.empty(int(L.b200mnist_conv2_bwd_scratch_bytes(B)), dtype=torch.uint8, device=dev)
        _cabi.check(L.b200mnist_conv2_bwd(dp2.data_ptr(), p2.data_ptr(), i2.data_ptr(), p1.data_ptr(),
                                          P["conv2.weight"].data_ptr(), B, G["conv2.weight"].data_ptr(),
                                          G["conv2.bias"].data_ptr(), dp1.data_ptr(), sc2.data_ptr(), s), "conv2 bwd")
        sc1 = torch.empty(int(L.b200mnist_conv1_bwd_scratch_bytes(B)), dtype=torch.uint8, device=dev)
        _cabi.check(L.b200mnist_conv1_bwd(dp1.data_ptr(), p1.data_ptr(), i1.data_ptr(), x.data_ptr(), B,
                                          G["conv1.weight"].data_ptr(), G["conv1.bias"].data_ptr(), sc1.data_ptr(), s),
                    "conv1 bwd")
        for name in NAMES:  # re-attach the views in case zero_grad() set them to None
            mod, attr = name.split(".")
            getattr(getattr(net, mod), attr).grad = G[name]
        if net._ddp is not None:
            net._ddp.reduce_gradients()
        return (None, None) + (None,) * 8


class StepGraph:
    """The fused training step captured once in a CUDA graph and replayed per batch
    (the B=64 step is launch-latency bound: ~10 kernels of a few microseconds each).
    Static input buffers ``x``/``target`` are filled by the caller (e.g. an H2D copy) before
    ``replay()``; ``loss`` is a 0-d view of the device scalar the graph writes.  With an ``InputRing`` a graph may
    hold ``unroll`` consecutive steps (each reads the slot under the ring's device cursor): one ``replay()`` is then
    ``unroll`` steps, and the ~1.5 us a graph boundary costs on the GPU is paid once per ``unroll`` steps."""

    def __init__(self, model: Net, optimizer, batch: int, x=None, target=None, loss_out=None, loss_host=None, ring=None,
                 unroll: int = 1):
        if optimizer.first_step:
            raise RuntimeError("run one eager step first (the first SGD step initialises the momentum buffer)")
        if unroll < 1 or (unroll > 1 and ring is None):
            raise ValueError("unroll > 1 needs an InputRing (each of the graph's steps reads its own slot)")
        self.unroll = unroll
        model = getattr(model, "module", model)  # accept our DistributedDataParallel wrapper
        dev = model.device
        self.model, self.optimizer = model, optimizer
        # static input buffers (the caller may supply its own, e.g. two views of one packed record)
        self.x = torch.zeros(batch, 1, 28, 28, dtype=torch.float32, device=dev) if x is None else x
        self.target = torch.zeros(batch, dtype=torch.int64, device=dev) if target is None else target
        model._check_inputs(self.x, self.target)
        self.lr, self.momentum = optimizer.lr, optimizer.momentum
        model.workspace(batch)  # allocate outside capture
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        # Warm-up on the capture stream (lazy CUDA/NCCL initialisation must happen outside the
        # capture); it is a real step on zero data, so the training state is put back afterwards.
        # With ``ring`` (an InputRing) every replay reads the ring slot under the ring's device cursor and advances
        # it: no input has to be staged between replays (x / target then receive a copy of the batch).
        self.ring = ring
        saved = [t.clone() for t in (model.flat_params, model.flat_grads, model.flat_momentum)]
        ring_snap = ring._snapshot() if ring is not None else None
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            model.train_step(self.x, self.target, optimizer, loss_out=loss_out, ring=ring)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize(dev)
        for t, sv in zip((model.flat_params, model.flat_grads, model.flat_momentum), saved):
            t.copy_(sv)
        if ring is not None:
            ring._restore(ring_snap)
        torch.cuda.synchronize(dev)
        # Under the fused DDP exchange a step writes its gradient into staging bucket (step & 1): the kernel
        # arguments differ between even and odd steps, so both variants are captured and replay() alternates.
        ddp = model._ddp
        self._ddp = ddp if (ddp is not None and ddp.fused) else None
        self.graphs = []
        for variant in range(2 if self._ddp is not None else 1):
            g = torch.cuda.CUDAGraph()
            if self._ddp is not None:
                self._ddp._force_parity = variant
            try:
                # thread_local: the NCCL watchdog thread may query its own events while this thread captures
                # (in the default global mode that query invalidates the capture and the ranks dead-lock)
                with torch.cuda.graph(g, stream=side, capture_error_mode="thread_local"):
                    for u in range(unroll):   # `unroll` consecutive steps per replay: one graph boundary per `unroll` steps
                        if self._ddp is not None:
                            self._ddp._force_parity = (variant + u) & 1
                        self.loss = model.train_step(self.x, self.target, optimizer, loss_out=loss_out, ring=ring)
                    if loss_host is not None:  # D2H memcpy node: every replay leaves the step's loss in pinned host memory
                        loss_host.copy_(self.loss.reshape(1), non_blocking=True)
            finally:
                if self._ddp is not None:
                    self._ddp._force_parity = None
            self.graphs.append(g)
        self.loss_host = loss_host
        model._graphs.append(weakref.ref(self))

    def release(self):
        """Destroy the captured graphs.  A graph that captured NCCL kernels must be gone before
        ``destroy_process_group()``, which otherwise blocks forever."""
        for g in self.graphs:
            g.reset()
        self.graphs = []

    def replay(self):
        if (self.optimizer.lr, self.optimizer.momentum) != (self.lr, self.momentum):
            raise RuntimeError("lr/momentum changed after capture; build a new StepGraph")
        if self._ddp is not None:
            variant = self._ddp._step & 1
            self._ddp._step += self.unroll
        else:
            variant = 0
        self.graphs[variant].replay()
        return self.loss


class StepLoss:
    """What ``StepPipeline.submit`` returns: the loss of the step just enqueued.  The step stores it
    in pinned host memory asynchronously, so reading it means waiting for THAT step -- ``item()`` (what
    mnist.py:47 calls) and ``float()`` wait until the step has reported completion to the host first, like
    ``loss.item()`` on a CUDA tensor does in the reference.  Read it before ``slots`` more steps are submitted."""
    __slots__ = ("_pipe", "_step")

    def __init__(self, pipe, step):
        self._pipe, self._step = pipe, step

    def item(self) -> float:
        self._pipe.wait_step(self._step)
        loss = self._pipe.ring.report(self._step)[1]
        if loss is None:
            raise RuntimeError("this step's loss has been overwritten: read it before `slots` more steps are submitted")
        return loss

    __float__ = item


class StepPipeline:
    """Host-fed training with nothing but graph launches on the compute stream.

    The batches travel through a small ring of device slots (``InputRing(host_fed=True)``): ``submit`` copies the
    pinned host batch of step n into slot n % slots on a COPY stream and writes the slot's sequence number behind the
    copy (a stream memory operation, ``b200mnist_stage_record``), then replays the step's CUDA graph.  The step's
    first kernel waits ON THE DEVICE until its slot carries its sequence number (normally it has for a long time: the
    copies run ahead of the steps) and its second-to-last kernel stores the loss and the step count into pinned host memory
    (one 8-byte word), which is how the host knows a slot may be refilled and a loss may be read -- no event, no stream
    wait, no copy node between two replays.

        pipe = StepPipeline(model, optimizer, batch)
        for rec in pinned_records:          # uint8 [batch*784*4 + batch*8]: fp32 images, then int64 targets
            loss = pipe.submit(rec)         # asynchronous; loss.item() waits for that step only
        pipe.drain()                        # pipe.loss_host[n % pipe.slots] holds the loss of step n (the last `slots` steps)
    """

    def __init__(self, model, optimizer, batch: int, slots: int = 8, unroll: int = 1):
        if unroll < 1 or 2 * unroll > slots:
            raise ValueError("unroll must be at least 1 and at most slots / 2 (the copies have to run ahead of the steps)")
        net = getattr(model, "module", model)
        dev = net.device
        self.batch, self.xbytes, self.rec_bytes, self.slots = batch, batch * 784 * 4, batch * 784 * 4 + batch * 8, slots
        row = (self.rec_bytes + 15) // 16 * 16
        self.records = torch.zeros(slots, row, dtype=torch.uint8, device=dev)
        self.ring = InputRing(batch, records=self.records, host_fed=True)
        self.x = torch.zeros(batch, 1, 28, 28, dtype=torch.float32, device=dev)     # the step's copy of its batch
        self.target = torch.zeros(batch, dtype=torch.int64, device=dev)
        self.graph = StepGraph(model, optimizer, batch, x=self.x, target=self.target, ring=self.ring)
        # `unroll` > 1: the steps are launched `unroll` at a time (one graph holding that many steps); a step somebody
        # waits for before its group is full goes out through the single-step graph
        self.unroll, self._staged = unroll, 0
        self.graph_n = (StepGraph(model, optimizer, batch, x=self.x, target=self.target, ring=self.ring, unroll=unroll)
                        if unroll > 1 else None)
        self.copy_stream = torch.cuda.Stream(device=dev)
        self._lib = _cabi.lib()
        self._keep = [None] * slots
        self.steps = 0

    @property
    def loss_host(self) -> torch.Tensor:
        """The losses last reported per ring slot (float32 [slots]; slot = step % slots)."""
        return (self.ring.report_host & 0xFFFFFFFF).to(torch.int32).view(torch.float32)

    def wait_step(self, step: int):
        """Block the host until step ``step`` (0-based) of this pipeline has reported to pinned host memory: its loss is
        final and its ring slot may be refilled (the report leaves with the step's second-to-last kernel; the SGD
        update completes right behind it, in stream order -- ``drain()`` or any stream-ordered read waits for that)."""
        if not 0 <= step < self.steps:
            raise IndexError(f"step {step} has not been submitted ({self.steps} so far)")
        if step >= self.steps - self._staged:
            self.flush()
        while not self.ring.report(step)[0]:
            pass

    def flush(self):
        """Launch the steps that were submitted but are still waiting for their group of ``unroll`` to fill (call it
        before anything else uses the model: another pipeline, an eager step, evaluation)."""
        while self._staged:
            self.graph.replay()
            self._staged -= 1

    def submit(self, host_record, target=None):
        """``host_record``: one packed uint8 record (fp32 images, then int64 targets), or the image tensor when
        ``target`` is given; pinned host memory (copied while earlier steps run) or device memory."""
        n = self.steps
        slot, seq = n % self.slots, n + 1
        if n >= self.slots:
            self.wait_step(n - self.slots)          # the step that last read this slot is done
        dst = self.records[slot].data_ptr()
        cs = self.copy_stream.cuda_stream
        ready = self.ring.landed.data_ptr()
        if host_record.is_cuda or (target is not None and target.is_cuda):
            # produced on the compute stream a moment ago (e.g. DeviceMNIST's gather kernel): the copy must follow it
            self.copy_stream.wait_stream(torch.cuda.current_stream())
        if target is None:
            if host_record.numel() * host_record.element_size() != self.rec_bytes or not host_record.is_contiguous():
                raise ValueError("a packed record is B*784*4 + B*8 contiguous bytes")
            _cabi.check(self._lib.b200mnist_stage_record(dst, host_record.data_ptr(), self.rec_bytes, cs, ready, seq),
                        "stage_record")
        else:
            if host_record.dtype != torch.float32 or target.dtype != torch.int64 or host_record.numel() != self.batch * 784 \
                    or target.numel() != self.batch or not host_record.is_contiguous() or not target.is_contiguous():
                raise ValueError("data must be contiguous float32 [B,1,28,28] and target int64 [B]")
            _cabi.check(self._lib.b200mnist_stage_record(dst, host_record.data_ptr(), self.xbytes, cs, 0, 0), "stage_record")
            _cabi.check(self._lib.b200mnist_stage_record(dst + self.xbytes, target.data_ptr(), self.batch * 8, cs, ready, seq),
                        "stage_record")
        self._keep[slot] = (host_record, target)        # the copy is asynchronous: keep the source alive until the slot is reused
        self.steps += 1
        self._staged += 1
        if self._staged == self.unroll:
            (self.graph_n or self.graph).replay()
            self._staged = 0
        elif self._device_starving():
            self.flush()        # nothing is running: grouping would only leave the GPU idle while the group fills
        return StepLoss(self, n)

    def steps_done(self) -> int:
        """Steps the device has completed on this pipeline's ring (reads the ring's device-side counter: synchronises)."""
        return int(self.ring.state[1])

    def _device_starving(self) -> bool:
        launched = self.steps - self._staged
        return launched == 0 or self.ring.report(launched - 1)[0]

    def drain(self):
        self.flush()
        torch.cuda.current_stream().synchronize()
        self.copy_stream.synchronize()
