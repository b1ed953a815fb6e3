"""Fused SGD for the training step: every parameter tensor updated by ONE kernel launch (asf_sgd_multi).

Same constructor arguments, parameter-group layout and update rule as ``torch.optim.SGD`` as the reference
configures it (/root/reference/audio_slowfast/models/optimizer.py:31-62: momentum 0.9, dampening 0, nesterov,
weight decay 1e-4 on non-BN parameters and 0 on the ``"bn"`` group; ``lr_policy.set_lr`` assigns
``param_group["lr"]`` every iteration).  The learning rate of each group lives in device memory and is refreshed
from a pinned host copy by an async memcpy at the start of ``step()``, so a step captured in a CUDA graph keeps
following the schedule.  ``construct_optimizer`` mirrors the reference helper of the same name.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib as L


class FusedSGD(torch.optim.Optimizer):
    def __init__(self, params, lr=0.1, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False):
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        super().__init__(params, dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                                      nesterov=nesterov))
        g0 = self.param_groups[0]
        for g in self.param_groups:
            if (g["momentum"], g["dampening"], g["nesterov"]) != (g0["momentum"], g0["dampening"], g0["nesterov"]):
                raise ValueError("FusedSGD: momentum / dampening / nesterov must be the same in every group")
        self._table = None
        self._table_key = None
        self._steps = 0
        self._lr_host = None
        self._lr_dev = None
        self._range_tables = {}
        self._range_done = None  # ids of the parameters already updated by step_range() in this iteration

    def _build_table(self, items, device):
        arr = (L.SgdDesc * len(items))()
        chunk = 0
        for i, (p, gi, wd, m) in enumerate(items):
            d = arr[i]
            d.p, d.g, d.m = p.data_ptr(), p.grad.data_ptr(), (m.data_ptr() if m is not None else None)
            d.n, d.wd, d.group, d.chunk0 = p.numel(), wd, gi, chunk
            chunk += -(-p.numel() // L.SGD_CHUNK)
        host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
        return host.to(device), len(items), chunk

    def refresh_lr(self):
        """Publishes the current ``param_group['lr']`` values to the pinned host copy (call before replaying a
        captured step; ``step()`` does it itself in eager mode)."""
        if self._lr_host is None:
            self._lr_host = torch.zeros(len(self.param_groups), dtype=torch.float32).pin_memory()
        for i, g in enumerate(self.param_groups):
            self._lr_host[i] = float(g["lr"])

    def _publish_lr(self, device):
        if self._lr_dev is None:
            self._lr_dev = torch.zeros(len(self.param_groups), dtype=torch.float32, device=device)
        if not torch.cuda.is_current_stream_capturing():
            self.refresh_lr()
        elif self._lr_host is None:
            self.refresh_lr()
        self._lr_dev.copy_(self._lr_host, non_blocking=True)

    @torch.no_grad()
    def begin_step(self, device):
        """Opens an iteration that will be updated piecewise by ``step_range``: publishes the learning rates on the
        current stream (every stream that later runs a piece must be ordered after this point)."""
        self._publish_lr(device)
        self._range_done = set()

    @torch.no_grad()
    def step_range(self, flat_grad, lo, hi):
        """Updates, on the current stream, the parameters whose gradients are the views of ``flat_grad`` lying inside
        elements [lo, hi) -- the piece of the engine's flat gradient buffer that has just become final -- so the
        optimiser runs under the rest of the backward pass.  ``step()`` afterwards updates whatever is left."""
        if self._range_done is None:
            raise RuntimeError("FusedSGD.step_range needs begin_step() first")
        base, esz = flat_grad.data_ptr(), flat_grad.element_size()
        mom = self.param_groups[0]["momentum"]
        items = []
        for gi, g in enumerate(self.param_groups):
            for p in g["params"]:
                if p.grad is None or id(p) in self._range_done:
                    continue
                off = (p.grad.data_ptr() - base) // esz
                if not (lo <= off and off + p.numel() <= hi) or p.grad.data_ptr() < base:
                    continue
                m = None
                if mom != 0:
                    st = self.state[p]
                    if "momentum_buffer" not in st:
                        st["momentum_buffer"] = torch.zeros_like(p)
                    m = st["momentum_buffer"]
                items.append((p, gi, float(g["weight_decay"]), m))
        if not items:
            return
        key = (lo, hi) + tuple((p.data_ptr(), p.grad.data_ptr(), wd) for p, _, wd, _ in items)
        table = self._range_tables.get(key)
        if table is None:
            if len(self._range_tables) > 64:
                self._range_tables.clear()
            table = self._range_tables[key] = self._build_table(items, items[0][0].device)
        self._launch(table, items)
        self._range_done.update(id(p) for p, _, _, _ in items)

    def _launch(self, table, items):
        g0 = self.param_groups[0]
        # first step: momentum buffer := d, which equals momentum*0 + (1-dampening)*d only for dampening == 0
        first = int(self._steps == 0 and g0["dampening"] != 0)
        L.call("asf_sgd_multi", L.ptr(table[0]), table[1], table[2], L.ptr(self._lr_dev),
               C.c_float(g0["momentum"]), C.c_float(g0["dampening"]), int(bool(g0["nesterov"])), first, L.stream_ptr())
        # The kernel wrote through raw pointers: tell autograd's version counters, which is what the engine keys its
        # packed bf16 weight copies on (Engine._prepare_all) -- without this the next forward would keep using the
        # weights from before the update.
        torch.autograd.graph.increment_version([p for p, _, _, _ in items])

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        items = []
        device = None
        mom = self.param_groups[0]["momentum"]
        done, self._range_done = self._range_done, None
        for gi, g in enumerate(self.param_groups):
            for p in g["params"]:
                if p.grad is None or (done is not None and id(p) in done):
                    continue
                if p.dtype != torch.float32 or p.grad.dtype != torch.float32 or not p.is_cuda:
                    raise RuntimeError("FusedSGD needs fp32 CUDA parameters and gradients")
                if not p.is_contiguous() or not p.grad.is_contiguous():
                    raise RuntimeError("FusedSGD needs contiguous parameters and gradients")
                m = None
                if mom != 0:
                    st = self.state[p]
                    if "momentum_buffer" not in st:
                        st["momentum_buffer"] = torch.zeros_like(p)
                    m = st["momentum_buffer"]
                items.append((p, gi, float(g["weight_decay"]), m))
                device = p.device
        if not items:
            if done:
                self._steps += 1
            return loss
        key = tuple((p.data_ptr(), p.grad.data_ptr(), wd) for p, _, wd, _ in items)
        if key != self._table_key:
            self._table = self._build_table(items, device)
            self._table_key = key
        if done is None:
            self._publish_lr(device)
        self._launch(self._table, items)
        self._steps += 1
        return loss


def construct_optimizer(model, cfg):
    """models/optimizer.py:11-71: BN parameters (``"bn" in name``) without weight decay, the rest with
    ``SOLVER.WEIGHT_DECAY``; SGD-nesterov per ``cfg.SOLVER`` -- on the fused kernel."""
    bn_params, non_bn = [], []
    for name, p in model.named_parameters():
        if not p.requires_grad:
            continue
        (bn_params if "bn" in name else non_bn).append(p)
    sol = cfg.SOLVER
    groups = [{"params": bn_params, "weight_decay": cfg.BN.WEIGHT_DECAY}, {"params": non_bn, "weight_decay": sol.WEIGHT_DECAY}]
    groups = [g for g in groups if g["params"]]
    if sol.OPTIMIZING_METHOD != "sgd":
        raise NotImplementedError("asf_b200.optim implements SOLVER.OPTIMIZING_METHOD 'sgd' (got %s)" % sol.OPTIMIZING_METHOD)
    return FusedSGD(groups, lr=sol.BASE_LR, momentum=sol.MOMENTUM, dampening=sol.DAMPENING,
                    weight_decay=sol.WEIGHT_DECAY, nesterov=sol.NESTEROV)
