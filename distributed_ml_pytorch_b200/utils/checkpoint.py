"""Checkpoint / resume in the plain ``torch`` ``state_dict`` format (NORTH-STAR; the reference has no
``torch.save`` / ``torch.load`` anywhere).

``model`` entries use exactly the key names and PyTorch shapes of the reference's modules
(example/models.py), so a file written here loads into the reference's ``LeNet`` / ``AlexNet`` with
``load_state_dict`` and vice versa; a bare ``torch.save(model.state_dict())`` file is accepted on load.
Optimizer state adds what the asynchronous optimizer needs to resume: step counter, the unpushed accumulator,
momentum, learning rate.
"""
from __future__ import annotations

from typing import Any, Dict

import torch


def model_state(model: torch.nn.Module) -> Dict[str, torch.Tensor]:
    return {k: v.detach().cpu().contiguous().clone() for k, v in model.state_dict().items()}


def optimizer_state(optimizer) -> Dict[str, Any]:
    st: Dict[str, Any] = {"class": type(optimizer).__name__, "lr": float(optimizer.param_groups[0]["lr"])}
    for name in ("idx", "n_push", "n_pull", "pushes", "pulls"):
        if hasattr(optimizer, name):
            st[name] = getattr(optimizer, name)
    acc = getattr(optimizer, "accumulated_gradients", None)
    if acc is not None:
        st["accumulated_gradients"] = acc.detach().cpu().clone()
    mom = getattr(optimizer, "momentum_buf", None)
    if mom is not None:
        st["momentum_buf"] = mom.detach().cpu().clone()
    return st


def save_checkpoint(path: str, model: torch.nn.Module, optimizer=None, **meta) -> None:
    payload = {"model": model_state(model), "meta": dict(meta), "format": "dmlb-checkpoint-v1"}
    if optimizer is not None:
        payload["optimizer"] = optimizer_state(optimizer)
    torch.save(payload, path)


def load_checkpoint(path: str, model: torch.nn.Module, optimizer=None, map_location="cpu") -> Dict[str, Any]:
    payload = torch.load(path, map_location=map_location, weights_only=False)
    if isinstance(payload, dict) and "model" in payload and "format" in payload:
        sd, meta, ost = payload["model"], payload.get("meta", {}), payload.get("optimizer")
    else:  # reference-style: the file *is* a state_dict
        sd, meta, ost = payload, {}, None
    model.load_state_dict(sd)
    arena = getattr(model, "_dmlb_arena", None)
    if arena is not None:
        arena.mark_params_changed()
    if optimizer is not None and ost is not None:
        optimizer.param_groups[0]["lr"] = ost.get("lr", optimizer.param_groups[0]["lr"])
        for name in ("idx", "pushes", "pulls"):
            if name in ost and hasattr(optimizer, name):
                setattr(optimizer, name, ost[name])
        if "accumulated_gradients" in ost and hasattr(optimizer, "accumulated_gradients"):
            n = min(ost["accumulated_gradients"].numel(), optimizer.accumulated_gradients.numel())
            optimizer.accumulated_gradients[:n].copy_(ost["accumulated_gradients"][:n])
        if "momentum_buf" in ost and getattr(optimizer, "momentum_buf", None) is not None:
            n = min(ost["momentum_buf"].numel(), optimizer.momentum_buf.numel())
            optimizer.momentum_buf[:n].copy_(ost["momentum_buf"][:n])
    return meta
