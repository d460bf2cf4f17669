"""Flat parameter / gradient arenas and the shard map of the parameter server.

The reference re-flattens the model on every optimizer step (``ravel_model_params`` allocates and
concatenates N floats, asgd/optim/Asynchronous.py:54) and updates parameters with one launch per tensor
(Asynchronous.py:63-68).  Here every parameter and every gradient of a model is a *view* into one
contiguous fp32 buffer, so

* ravel / unravel are zero-copy slices (``utils/serialization.py``);
* the optimizer, the gradient push, the parameter pull and the synchronous reduce-scatter/all-gather are
  single kernels over ``[0, N)``;
* weight-gradient GEMMs accumulate straight into the arena (``wsink`` / ``bsink`` of the nn layers).

:class:`ShardMap` splits ``[0, N)`` into contiguous, 1 KiB-aligned shards, one per owning GPU
(SURVEY.md §5.1 "sharded PS"); the reference has a single server rank (example/main.py:151,166-167).
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch
import torch.nn as nn

ALIGN_ELEMS = 256  # 1 KiB of fp32: keeps every shard boundary 16-byte (float4) and sector aligned


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class ShardMap:
    """Contiguous partition of a flat vector of ``numel`` elements over ``n_shards`` owners."""

    def __init__(self, numel: int, n_shards: int, align: int = ALIGN_ELEMS):
        if n_shards < 1:
            raise ValueError("n_shards must be >= 1")
        self.numel = int(numel)
        self.n_shards = int(n_shards)
        self.shard_size = _round_up(max(1, -(-self.numel // self.n_shards)), align)
        self.padded = self.shard_size * self.n_shards

    def owner_of(self, index: int) -> int:
        if not 0 <= index < self.padded:
            raise IndexError(index)
        return index // self.shard_size

    def bounds(self, shard: int) -> Tuple[int, int]:
        """[begin, end) of ``shard`` in the padded flat vector."""
        if not 0 <= shard < self.n_shards:
            raise IndexError(shard)
        return shard * self.shard_size, (shard + 1) * self.shard_size

    def valid_bounds(self, shard: int) -> Tuple[int, int]:
        """[begin, end) clipped to the real (unpadded) length."""
        b, e = self.bounds(shard)
        return min(b, self.numel), min(e, self.numel)

    def __repr__(self):
        return f"ShardMap(numel={self.numel}, n_shards={self.n_shards}, shard_size={self.shard_size})"


class FlatArena:
    """Owns the flat fp32 ``params`` and ``grads`` buffers of one model replica and re-points every
    ``Parameter`` (and its ``.grad``) at views of them.

    ``params_storage`` / ``grads_storage`` may be supplied by the caller — the multi-GPU paths pass tensors
    that wrap NVLink-peer-mapped allocations so other GPUs can read / write them directly.
    """

    def __init__(self, model: nn.Module, n_shards: int = 1, params_storage: Optional[torch.Tensor] = None,
                 grads_storage: Optional[torch.Tensor] = None, align: int = ALIGN_ELEMS):
        self.model = model
        plist = [p for p in model.parameters()]
        if not plist:
            raise ValueError("model has no parameters")
        self.device = plist[0].device
        self.numel = sum(p.numel() for p in plist)
        self.shards = ShardMap(self.numel, n_shards, align=align)
        self.padded = self.shards.padded
        self.params = params_storage if params_storage is not None else torch.zeros(self.padded, dtype=torch.float32, device=self.device)
        self.grads = grads_storage if grads_storage is not None else torch.zeros(self.padded, dtype=torch.float32, device=self.device)
        if self.params.numel() < self.padded or self.grads.numel() < self.padded:
            raise ValueError("supplied arena storage is too small")
        self.offsets: Dict[str, Tuple[int, int]] = {}
        self._bound = False
        self.channels_last = self.params.is_cuda
        # flat bf16 mirror of the arena (same offsets): the optimizer kernel writes it alongside the fp32 masters.  A tf32 model reads
        # the fp32 masters themselves and keeps no mirror.
        self.compute_dtype = getattr(model, "_dmlb_compute_dtype", torch.bfloat16)
        want_mirror = self.params.is_cuda and self.compute_dtype == torch.bfloat16
        self.shadow = torch.zeros(self.padded, dtype=torch.bfloat16, device=self.device) if want_mirror else None
        self.bind()

    def is_bound(self) -> bool:
        return self._bound

    def _view(self, flat: torch.Tensor, off: int, p: torch.Tensor) -> torch.Tensor:
        """View of ``flat[off:off+numel]`` with ``p``'s logical shape.  On CUDA, conv weights are stored
        channels-last (physically ``[Cout, kh, kw, Cin]`` — exactly the K order of the implicit GEMM), so their bf16
        shadows are plain casts and weight-gradient GEMMs write contiguous rows; the tensor the user sees keeps
        PyTorch's logical ``[Cout, Cin, kh, kw]`` shape, like ``memory_format=torch.channels_last`` weights."""
        n = p.numel()
        if self.channels_last and p.dim() == 4:
            co, ci, kh, kw = p.shape
            return flat[off:off + n].view(co, kh, kw, ci).permute(0, 3, 1, 2)
        return flat[off:off + n].view(p.shape)

    def bind(self) -> None:
        off = 0
        with torch.no_grad():
            for name, p in self.model.named_parameters():
                n = p.numel()
                view = self._view(self.params, off, p)
                view.copy_(p.detach().to(self.params.device))
                p.data = view
                p.grad = self._view(self.grads, off, p)
                self.offsets[name] = (off, n)
                off += n
            if off < self.padded:
                self.params[off:].zero_()
            self.grads.zero_()
        self._attach_sinks()
        self.model._dmlb_arena = self
        self._bound = True
        shadows = getattr(self.model, "_dmlb_shadows", None)
        if shadows is not None:
            shadows.invalidate_table()
        self.mark_params_changed()

    def _attach_sinks(self) -> None:
        # hand the gradient views (and flat-shadow slices) to kernel-backed layers so wgrad GEMMs accumulate in place
        by_param = {id(p): name for name, p in self.model.named_parameters()}
        for mod in self.model.modules():
            if hasattr(mod, "_wsink") and getattr(mod, "weight", None) is not None:
                mod._wsink = mod.weight.grad
                mod._bsink = mod.bias.grad if getattr(mod, "bias", None) is not None else None
                name = by_param.get(id(mod.weight))
                mod._arena_slot = (self, self.offsets[name][0]) if (name is not None and self.params.is_cuda) else None

    def rebind_grads(self) -> None:
        """Re-attach ``.grad`` views (after a foreign ``zero_grad(set_to_none=True)``)."""
        for name, p in self.model.named_parameters():
            off, n = self.offsets[name]
            p.grad = self._view(self.grads, off, p)
        self._attach_sinks()

    def zero_grads(self) -> None:
        self.grads.zero_()

    def mark_params_changed(self, flat_done: bool = False, table_done: bool = False) -> None:
        """Call after writing ``params`` through raw pointers: refreshes the weight shadows (``flat_done``: the writing kernel
        already updated the flat bf16 mirror ``self.shadow``; ``table_done``: it also refreshed the table-backed shadows)."""
        shadows = getattr(self.model, "_dmlb_shadows", None)
        if shadows is not None and self.params.is_cuda:
            shadows.refresh(flat_done=flat_done, table_done=table_done)

    def fused_shadow_table(self):
        """Argument for ``K.sgd_flat_ex(table=...)``: the model's table-backed shadows, refreshed inside the optimizer kernel."""
        shadows = getattr(self.model, "_dmlb_shadows", None)
        if shadows is None or not self.params.is_cuda:
            return None
        return shadows.fused_table(self)

    def flat_params(self) -> torch.Tensor:
        return self.params[: self.numel]

    def flat_grads(self) -> torch.Tensor:
        return self.grads[: self.numel]


def ensure_arena(model: nn.Module, n_shards: int = 1, **kw) -> FlatArena:
    """Return the model's arena, creating it when absent or when the model moved to another device."""
    arena = getattr(model, "_dmlb_arena", None)
    dev = next(model.parameters()).device
    if arena is None or arena.device != dev or arena.shards.n_shards != n_shards or kw:
        arena = FlatArena(model, n_shards=n_shards, **kw)
    return arena
