"""Synchronous data-parallel SGD (NORTH-STAR; the reference has no synchronous mode).

Baseline formulation: ``all_reduce(grads) / world`` followed by ``optim.SGD.step()``.  Here, on GPUs, one kernel per
step and GPU (``sync_rs_update_ag``, csrc/comm.cu K4) does barrier -> reduce-scatter of the gradient arena over
NVLink peer loads -> SGD / momentum on the owner's fp32 shard -> all-gather of the new weights by peer stores into
every replica -> barrier.  The parameter and gradient arenas of every replica live in CUDA-IPC-shared memory for
that purpose.  On CPU tensors (gloo) the same optimizer falls back to ``dist.all_reduce`` + an in-place update.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

from .arena import FlatArena
from .fused_sgd import ArenaOptimizer


class Synchronous(ArenaOptimizer):
    def __init__(self, params, lr: float, model: torch.nn.Module, momentum: float = 0.0, weight_decay: float = 0.0, group=None):
        if lr < 0.0:
            raise ValueError("Invalid learning rate: {}".format(lr))
        device = next(model.parameters()).device
        self.group = group
        self.is_peer = device.type == "cuda"
        arena = None
        if self.is_peer:
            from . import peer
            self.peer = peer.PeerGroup(group=group, device=device)
            n = sum(p.numel() for p in model.parameters())
            world = self.peer.world
            shard = ((n + world - 1) // world + 255) // 256 * 256
            padded = shard * world
            self.params_sym = self.peer.symmetric(padded * 4)
            self.grads_sym = self.peer.symmetric(padded * 4)
            self.flags_sym = self.peer.symmetric(peer.flag_words() * 4)
            arena = FlatArena(model, n_shards=world, params_storage=self.params_sym.view(torch.float32, padded),
                              grads_storage=self.grads_sym.view(torch.float32, padded))
            assert arena.shards.shard_size == shard
            self.world = world
        else:
            self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        super().__init__(params, dict(lr=lr, momentum=momentum, weight_decay=weight_decay), model, arena=arena)
        a = self.arena
        if self.is_peer:
            self.momentum_buf = torch.zeros(a.shards.shard_size, dtype=torch.float32, device=device) if momentum else None
            self.peer.barrier()
        else:
            self.momentum_buf = torch.zeros_like(a.params) if momentum else None
        self.broadcast_parameters()

    def broadcast_parameters(self, src: int = 0) -> None:
        """Make every replica start from rank ``src``'s parameters."""
        if self.world == 1 or not (dist.is_available() and dist.is_initialized()):
            return
        dist.broadcast(self.arena.params, src=src, group=self.group)
        self.arena.mark_params_changed()
        if self.is_peer:
            self.peer.barrier()

    def step_local(self) -> None:
        """Nothing is local in synchronous mode: the whole update is the fused collective in :meth:`communicate`."""
        return None

    @torch.no_grad()
    def communicate(self) -> None:
        g = self.param_groups[0]
        a = self.arena
        self._sync_lr()
        if self.is_peer:
            from . import peer
            peer.sync_rs_update_ag(self.grads_sym.table, self.params_sym.table, self.flags_sym.table, self.momentum_buf, self.peer.rank,
                                   self.world, a.shards.shard_size, self.lr_dev, g["momentum"], g["weight_decay"], 1.0 / self.world)
            a.mark_params_changed()
        else:
            if self.world > 1:
                dist.all_reduce(a.grads, group=self.group)
                a.grads.div_(self.world)
            d = a.grads
            if g["weight_decay"]:
                d = d.add(a.params, alpha=g["weight_decay"])
            if self.momentum_buf is not None:
                self.momentum_buf.mul_(g["momentum"]).add_(d)
                d = self.momentum_buf
            a.params.add_(d, alpha=-g["lr"])

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        self.communicate()
        return loss

    def finish(self) -> None:
        if self.is_peer:
            self.peer.barrier()
