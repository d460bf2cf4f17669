"""Model zoo: the reference's LeNet and AlexNet (example/models.py) plus the NORTH-STAR ResNet-18/50."""
from __future__ import annotations

from .alexnet import AlexNet
from .lenet import LeNet

__all__ = ["LeNet", "AlexNet", "build_model", "default_dtype", "MODEL_NAMES"]

MODEL_NAMES = ("lenet", "alexnet", "resnet18", "resnet50")


def default_dtype(name: str) -> str:
    """CUDA compute dtype a model runs in unless told otherwise: the reference's own models at the reference's precision
    (fp32 tensors, TF32 tensor-core math — example/main.py:41,53-54,70-73), the NORTH-STAR ResNets in bf16 (BASELINE.json config 4)."""
    return "tf32" if name.lower() in ("lenet", "alexnet") else "bf16"


def build_model(name: str, num_classes: int = 10, in_channels: int = 3, input_hw: int = 32, dtype=None):
    """Factory used by ``example/main.py --model`` (the reference hard-codes AlexNet, main.py:41,136).
    ``dtype``: ``'tf32'`` | ``'bf16'`` | ``None`` (= :func:`default_dtype`); it only matters on CUDA."""
    from ..nn import set_compute_dtype
    name = name.lower()
    if name == "lenet":
        model = LeNet(in_channels=in_channels, input_hw=input_hw, num_classes=num_classes, dtype=dtype or "tf32")
    elif name == "alexnet":
        model = AlexNet(num_classes=num_classes, in_channels=in_channels, dtype=dtype or "tf32")
    elif name in ("resnet18", "resnet50"):
        from . import resnet
        model = getattr(resnet, name)(num_classes=num_classes, in_channels=in_channels, small_input=input_hw <= 64)
    else:
        raise ValueError(f"unknown model {name!r}; choose from {MODEL_NAMES}")
    return set_compute_dtype(model, dtype if dtype is not None else default_dtype(name))
