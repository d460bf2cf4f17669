"""Flat (de)serialisation of a model's parameters.

Re-specifies the reference's ABSENT ``asgd.utils.serialization`` from its call sites
(reference asgd/optim/Asynchronous.py:18,27,34,54):

* ``ravel_model_params(model, grads=False)`` -> 1-D tensor of every parameter (or every ``.grad``)
  in ``model.parameters()`` order;
* ``unravel_model_params(model, flat)`` -> in-place copy of a flat tensor back into the parameters.

B200-first twist: when the model has been placed in a :class:`~distributed_ml_pytorch_b200.parallel.arena.FlatArena`
every parameter (and gradient) is already a view into one contiguous buffer, so ravel is a zero-copy
slice and unravel is a single ``copy_`` — the per-step ``cat`` + allocation of the reference design
(SURVEY.md §2.5 "ravel_model_params ... every step") disappears.  (CUDA arenas store conv weights channels-last
physically; there these two functions keep returning the *logical* ``model.parameters()`` order — which costs a
gather — while the optimizers and the push / pull / sync kernels work on the arena's own flat order directly.)
"""
from __future__ import annotations

from typing import List

import torch


def _params(model) -> List[torch.nn.Parameter]:
    if isinstance(model, torch.nn.Module):
        return list(model.parameters())
    return list(model)


def numel(model) -> int:
    """Total number of scalar parameters (``N`` in SURVEY.md)."""
    return sum(p.numel() for p in _params(model))


def ravel_model_params(model, grads: bool = False) -> torch.Tensor:
    """Return all parameters (``grads=False``) or all gradients (``grads=True``) as one 1-D tensor.

    Missing gradients are raveled as zeros so the result always has length ``N``.
    """
    arena = getattr(model, "_dmlb_arena", None)
    if arena is not None and arena.is_bound() and not arena.channels_last:
        flat = arena.grads if grads else arena.params
        return flat[: arena.numel]
    parts = []
    for p in _params(model):
        if grads:
            g = p.grad
            parts.append(torch.zeros_like(p).reshape(-1) if g is None else g.detach().reshape(-1))
        else:
            parts.append(p.detach().reshape(-1))
    if not parts:
        return torch.zeros(0)
    return torch.cat(parts)


def unravel_model_params(model, flat: torch.Tensor) -> None:
    """Copy ``flat`` (length >= N) into the model's parameters in place."""
    arena = getattr(model, "_dmlb_arena", None)
    if arena is not None and arena.is_bound() and not arena.channels_last:
        with torch.no_grad():
            arena.params[: arena.numel].copy_(flat[: arena.numel].to(arena.params.device, non_blocking=True))
        arena.mark_params_changed()
        return
    offset = 0
    with torch.no_grad():
        for p in _params(model):
            n = p.numel()
            p.copy_(flat[offset:offset + n].view(p.shape))
            offset += n
    if arena is not None and arena.is_bound():
        arena.mark_params_changed()


def unravel_model_grads(model, flat: torch.Tensor) -> None:
    """Copy ``flat`` into the parameters' ``.grad`` fields (allocating them when absent)."""
    offset = 0
    for p in _params(model):
        n = p.numel()
        g = flat[offset:offset + n].view_as(p)
        if p.grad is None:
            p.grad = g.clone()
        else:
            p.grad.copy_(g)
        offset += n
