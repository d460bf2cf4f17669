"""Minimal stand-in for the reference's ABSENT asgd.utils.serialization (contract: SURVEY.md §0.2 A1).

The reference optimizer keeps its accumulator on the CPU (asgd/optim/Asynchronous.py:27) and adds the
raveled gradients to it (:55), so the raveled vector is returned as a CPU tensor even for a CUDA model.
"""
import torch


def ravel_model_params(model, grads=False):
    parts = []
    for p in model.parameters():
        t = p.grad if grads else p.data
        if t is None:
            t = torch.zeros_like(p.data)
        parts.append(t.detach().reshape(-1).cpu())
    return torch.cat(parts)


def unravel_model_params(model, flat):
    off = 0
    for p in model.parameters():
        n = p.numel()
        p.data.copy_(flat[off:off + n].view_as(p.data))
        off += n
