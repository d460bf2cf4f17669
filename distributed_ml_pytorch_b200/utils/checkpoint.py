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


def _arena_of(optimizer):
    return getattr(optimizer, "arena", None)


def _to_logical(arena, buf: torch.Tensor) -> torch.Tensor:
    """Flat optimizer buffer (arena order: conv weights channels-last on CUDA, padded to the shard grid) -> ``model.parameters()``
    order with PyTorch's element order, length N.  That is the order ``ravel_model_params`` defines, so a checkpoint does not depend
    on the device, the world size or the chunk alignment it was written with."""
    if arena is None:
        return buf.detach().cpu().clone()
    parts = []
    for name, p in arena.model.named_parameters():
        off, _ = arena.offsets[name]
        parts.append(arena._view(buf, off, p).detach().reshape(-1).cpu())
    return torch.cat(parts)


def _from_logical(arena, vec: torch.Tensor, buf: torch.Tensor) -> None:
    if arena is None:
        buf[: vec.numel()].copy_(vec)
        return
    o = 0
    for name, p in arena.model.named_parameters():
        off, n = arena.offsets[name]
        arena._view(buf, off, p).copy_(vec[o:o + n].view(p.shape))
        o += n


def server_state_dict(model: torch.nn.Module, optimizer) -> Dict[str, torch.Tensor]:
    """``state_dict`` of the *server's* parameters for an asynchronous run on the NVLink transport (the replica a worker holds lags
    the sharded masters by up to ``n_pull`` steps); any other optimizer: the model's own ``state_dict``."""
    sd = model_state(model)
    if getattr(optimizer, "transport", None) != "peer":
        return sd
    arena = optimizer.arena
    flat = optimizer.server_parameters()
    for name, p in model.named_parameters():
        off, _ = arena.offsets[name]
        sd[name] = arena._view(flat, off, p).detach().cpu().contiguous().clone()
    return sd


def optimizer_state(optimizer) -> Dict[str, Any]:
    st: Dict[str, Any] = {"class": type(optimizer).__name__, "lr": float(optimizer.param_groups[0]["lr"]), "layout": "logical"}
    for name in ("idx", "n_push", "n_pull", "pushes", "pulls"):
        if hasattr(optimizer, name):
            st[name] = getattr(optimizer, name)
    arena = _arena_of(optimizer)
    acc = getattr(optimizer, "accumulated_gradients", None)
    if acc is not None:
        st["accumulated_gradients"] = _to_logical(arena, acc)
    mom = getattr(optimizer, "momentum_buf", None)
    if mom is not None:
        st["momentum_buf"] = _to_logical(arena, mom)
    return st


def save_checkpoint(path: str, model: torch.nn.Module, optimizer=None, **meta) -> None:
    """Write ``path`` atomically (temporary file + rename).  With an asynchronous optimizer on the NVLink transport the model entries
    are the server's parameters (:func:`server_state_dict`)."""
    import os
    payload = {"model": server_state_dict(model, optimizer) if optimizer is not None else model_state(model), "meta": dict(meta),
               "format": "dmlb-checkpoint-v2"}
    if optimizer is not None:
        payload["optimizer"] = optimizer_state(optimizer)
    tmp = f"{path}.tmp{os.getpid()}"
    torch.save(payload, tmp)
    os.replace(tmp, path)


def read_checkpoint(path: str, map_location="cpu"):
    """-> (model state_dict, meta, optimizer state or None); accepts this package's files and a bare ``state_dict``."""
    payload = torch.load(path, map_location=map_location, weights_only=False)
    if isinstance(payload, dict) and "model" in payload and "format" in payload:
        return payload["model"], payload.get("meta", {}), payload.get("optimizer")
    return payload, {}, None   # reference-style: the file *is* a state_dict


def load_model_state(model: torch.nn.Module, sd) -> None:
    model.load_state_dict(sd)
    arena = getattr(model, "_dmlb_arena", None)
    if arena is not None:
        arena.mark_params_changed()


def load_optimizer_state(optimizer, ost) -> None:
    if optimizer is None or ost is None:
        return
    optimizer.param_groups[0]["lr"] = ost.get("lr", optimizer.param_groups[0]["lr"])
    for name in ("idx", "pushes", "pulls"):
        if name in ost and hasattr(optimizer, name):
            setattr(optimizer, name, ost[name])
    arena = _arena_of(optimizer)
    for key in ("accumulated_gradients", "momentum_buf"):
        buf = getattr(optimizer, key, None)
        if key not in ost or buf is None:
            continue
        vec = ost[key]
        if ost.get("layout") == "logical":
            want = arena.numel if arena is not None else buf.numel()
            if vec.numel() != want:
                raise ValueError(f"checkpoint {key} has {vec.numel()} elements, the model has {want} parameters")
            _from_logical(arena, vec.to(buf.device), buf)
        else:
            # v1 files stored the raw arena (device layout, padded to the writer's shard grid): only a same-layout reader may take it
            if arena is not None and (arena.channels_last or vec.numel() != buf.numel()):
                raise ValueError(f"checkpoint {key} was written in a device-specific arena layout (format v1); it cannot be restored "
                                 "into this arena — re-save it with this version or drop the optimizer state")
            buf[: vec.numel()].copy_(vec)
        if key == "accumulated_gradients" and hasattr(optimizer, "set_unpushed"):
            optimizer.set_unpushed(buf)       # overlap mode keeps no accumulator: the snapshot is rewound instead


def load_checkpoint(path: str, model: torch.nn.Module, optimizer=None, map_location="cpu") -> Dict[str, Any]:
    """Restore model (+ optimizer) from ``path``.  A distributed asynchronous optimizer seeds the parameter server when it is
    *constructed*, so for a resume load the model first (``read_checkpoint`` + ``load_model_state``), build the optimizer, then
    ``load_optimizer_state`` — ``example/main.py --resume`` does exactly that; this one-call form is for local optimizers."""
    sd, meta, ost = read_checkpoint(path, map_location)
    load_model_state(model, sd)
    load_optimizer_state(optimizer, ost)
    return meta
