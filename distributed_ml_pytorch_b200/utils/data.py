"""Input pipelines.

* :class:`GpuResidentLoader` — the whole (uint8) image set lives in HBM; every batch is one kernel launch that gathers a shuffled
  index set, converts HWC uint8 -> CHW fp32 and normalises (``dmlb_gather_normalize_u8``).  This is the B200-first replacement for
  the reference's host pipeline (torchvision ``ToTensor`` + ``Normalize`` in a 1-worker ``DataLoader`` and a pageable ``.cuda()``
  copy per batch, reference example/main.py:23-29,35-40,64-65), which cannot feed a step that consumes 3x10^5 images/s.
  On CPU tensors the same class computes the identical transform with torch ops.
* :func:`cifar10_arrays` — reads an already-downloaded torchvision CIFAR-10 directory into uint8 arrays (no network here).
"""
from __future__ import annotations

import os
import pickle
from typing import Iterator, Optional, Tuple

import numpy as np
import torch


class GpuResidentLoader:
    """Iterate shuffled, normalised batches from an image set resident on ``device``.

    ``images``: uint8 ``[N, H, W, C]`` (torchvision's ``CIFAR10.data`` layout); ``labels``: int ``[N]``.
    Yields ``(x, y)`` with ``x`` fp32 ``[B, C, H, W]`` = ``(images/255 - mean)/std`` and ``y`` int64 ``[B]`` — exactly
    ``Normalize(mean, std)(ToTensor(img))`` of the reference.  ``drop_last`` keeps the batch shape fixed for the CUDA-graph step.
    """

    def __init__(self, images, labels, batch_size: int, device, shuffle: bool = True, drop_last: bool = True, mean=(0.5, 0.5, 0.5),
                 std=(0.5, 0.5, 0.5), seed: int = 0):
        self.device = torch.device(device)
        self.images = torch.as_tensor(np.ascontiguousarray(images) if isinstance(images, np.ndarray) else images).to(self.device).contiguous()
        if self.images.dtype != torch.uint8 or self.images.dim() != 4:
            raise ValueError("images must be uint8 [N, H, W, C]")
        self.labels = torch.as_tensor(np.asarray(labels) if not torch.is_tensor(labels) else labels).to(self.device, dtype=torch.int64).contiguous()
        self.batch_size, self.shuffle, self.drop_last = int(batch_size), shuffle, drop_last
        self.mean, self.std = tuple(float(m) for m in mean), tuple(float(s) for s in std)
        self._gen = torch.Generator(device=self.device)
        self._gen.manual_seed(seed)
        import ctypes
        self._mean_c = (ctypes.c_float * 3)(*self.mean)
        self._std_c = (ctypes.c_float * 3)(*self.std)

    def __len__(self) -> int:
        n = self.images.shape[0]
        return n // self.batch_size if self.drop_last else -(-n // self.batch_size)

    def gather(self, index: torch.Tensor, out_x: Optional[torch.Tensor] = None, out_y: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        """Batch for the given sample indices (optionally written straight into caller-provided buffers, e.g. TrainStep.static_x)."""
        N, H, W, C = self.images.shape
        B = index.numel()
        if out_x is None:
            out_x = torch.empty((B, C, H, W), dtype=torch.float32, device=self.device)
        if out_y is None:
            out_y = torch.empty((B,), dtype=torch.int64, device=self.device)
        if self.device.type == "cuda":
            from .. import _native
            _native.call("dmlb_gather_normalize_u8", self.images.data_ptr(), self.labels.data_ptr(), index.data_ptr(), out_x.data_ptr(),
                         out_y.data_ptr(), B, H, W, C, self._mean_c, self._std_c, torch.cuda.current_stream(self.device).cuda_stream)
        else:
            x = self.images[index].permute(0, 3, 1, 2).float().div_(255.0)
            m = torch.tensor((self.mean + self.mean[-1:] * C)[:C]).view(1, C, 1, 1)
            s = torch.tensor((self.std + self.std[-1:] * C)[:C]).view(1, C, 1, 1)
            out_x.copy_((x - m) / s)
            out_y.copy_(self.labels[index])
        return out_x, out_y

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        n = self.images.shape[0]
        order = torch.randperm(n, device=self.device, generator=self._gen) if self.shuffle else torch.arange(n, device=self.device)
        for i in range(len(self)):
            idx = order[i * self.batch_size:(i + 1) * self.batch_size]
            yield self.gather(idx)


def cifar10_arrays(root: str, train: bool = True):
    """uint8 images ``[N, 32, 32, 3]`` and labels from a torchvision-style ``cifar-10-batches-py`` directory under ``root``."""
    base = os.path.join(root, "cifar-10-batches-py")
    names = [f"data_batch_{i}" for i in range(1, 6)] if train else ["test_batch"]
    xs, ys = [], []
    for n in names:
        with open(os.path.join(base, n), "rb") as f:
            d = pickle.load(f, encoding="latin1")
        xs.append(np.asarray(d["data"], dtype=np.uint8).reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1))
        ys.extend(d["labels"] if "labels" in d else d["fine_labels"])
    return np.ascontiguousarray(np.concatenate(xs)), np.asarray(ys, dtype=np.int64)
