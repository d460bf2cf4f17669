"""Arena-backed optimizers: one kernel launch updates the whole model.

:class:`FusedSGD` is the ``--no-distributed`` optimizer (the reference uses ``optim.SGD(lr, momentum=0.0)``,
example/main.py:43-44); :class:`ArenaOptimizer` is the base that the asynchronous and synchronous
distributed optimizers build on.  All of them stay ``torch.optim.Optimizer`` subclasses so LR schedulers
(reference main.py:47-48) and ``state_dict()`` keep working.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch.optim.optimizer import Optimizer

from .arena import FlatArena, ensure_arena


class ArenaOptimizer(Optimizer):
    """Base class: binds the model to a :class:`FlatArena` and keeps the learning rate in device memory
    (kernels read it through a pointer, so a captured CUDA graph follows LR-scheduler updates)."""

    def __init__(self, params, defaults, model: torch.nn.Module, n_shards: int = 1, arena: Optional[FlatArena] = None):
        if model is None:
            raise ValueError("model is required (the optimizer works on the model's flat arena)")
        super().__init__(params, defaults)
        self.model = model
        self.arena = arena if arena is not None else ensure_arena(model, n_shards=n_shards)
        self.device = self.arena.params.device
        self._lr_cached = None
        # When True the update kernel clears each gradient as it consumes it, and zero_grad() becomes a no-op
        # (TrainStep turns this on: one memset node less per step).  After step() the .grad views then read zero.
        self.fused_zero_grad = False
        self.lr_dev = torch.zeros(1, dtype=torch.float32, device=self.device)
        self._sync_lr()

    # -- learning rate -------------------------------------------------------------------------
    @property
    def lr(self) -> float:
        return float(self.param_groups[0]["lr"])  # single group, like the reference (Asynchronous.py:52)

    def _sync_lr(self) -> None:
        lr = self.lr
        if lr != self._lr_cached:
            self.lr_dev.fill_(lr)
            self._lr_cached = lr

    # -- gradients -----------------------------------------------------------------------------
    def zero_grad(self, set_to_none: bool = False) -> None:
        """One memset of the gradient arena (gradients stay views, never ``None``)."""
        first = next(iter(self.model.parameters()))
        self._check_model_still_bound(first)
        if self.fused_zero_grad and self.arena.params.is_cuda:
            return
        if first.grad is None or first.grad.data_ptr() != self.arena.grads.data_ptr():
            self.arena.rebind_grads()
        self.arena.zero_grads()

    def _check_model_still_bound(self, first) -> None:
        """The reference builds its optimizer and only then calls ``net.cuda()`` (example/main.py:46,53-54).  With a flat arena
        that would silently train a detached copy, so it is reported instead (build the optimizer after the move)."""
        if first.data_ptr() != self.arena.params.data_ptr():
            if first.device != self.arena.params.device:
                raise RuntimeError(
                    f"the model moved to {first.device} after the optimizer was created on {self.arena.params.device}: "
                    "construct the optimizer after model.to(device) / model.cuda()")
            self.arena.bind()          # same device, parameters re-created (e.g. load_state_dict(assign=True)): re-attach them

    def refresh_shadows(self, flat_done: bool = False) -> None:
        self.arena.mark_params_changed(flat_done=flat_done)


class FusedSGD(ArenaOptimizer):
    """SGD (+momentum, +weight decay) over the flat arena: ``theta -= lr * (mu * v + g)`` in one launch."""

    def __init__(self, params, lr: float, momentum: float = 0.0, weight_decay: float = 0.0, model: torch.nn.Module = None):
        if lr < 0.0:
            raise ValueError(f"Invalid learning rate: {lr}")
        super().__init__(params, dict(lr=lr, momentum=momentum, weight_decay=weight_decay), model)
        self.momentum_buf = torch.zeros_like(self.arena.params) if momentum != 0.0 else None

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        g = self.param_groups[0]
        self._sync_lr()
        a = self.arena
        if a.params.is_cuda:
            from ..ops import kernels as K
            table = a.fused_shadow_table()
            K.sgd_flat_ex(a.params, a.grads, None, self.momentum_buf, a.padded, self.lr_dev, g["momentum"], g["weight_decay"],
                          shadow=a.shadow, zero_grads=self.fused_zero_grad, table=table)
            a.mark_params_changed(flat_done=a.shadow is not None, table_done=table is not None)
        else:
            d = a.grads
            if g["weight_decay"]:
                d = d.add(a.params, alpha=g["weight_decay"])
            if self.momentum_buf is not None:
                self.momentum_buf.mul_(g["momentum"]).add_(d)
                d = self.momentum_buf
            a.params.add_(d, alpha=-g["lr"])
        return loss
