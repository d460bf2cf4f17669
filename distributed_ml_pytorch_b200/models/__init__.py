"""Model zoo: the reference's LeNet and AlexNet (example/models.py) plus the NORTH-STAR ResNet-18/50."""
from __future__ import annotations

from .alexnet import AlexNet
from .lenet import LeNet

__all__ = ["LeNet", "AlexNet", "build_model", "MODEL_NAMES"]

MODEL_NAMES = ("lenet", "alexnet", "resnet18", "resnet50")


def build_model(name: str, num_classes: int = 10, in_channels: int = 3, input_hw: int = 32):
    """Factory used by ``example/main.py --model`` (the reference hard-codes AlexNet, main.py:41,136)."""
    name = name.lower()
    if name == "lenet":
        return LeNet(in_channels=in_channels, input_hw=input_hw, num_classes=num_classes)
    if name == "alexnet":
        return AlexNet(num_classes=num_classes, in_channels=in_channels)
    if name in ("resnet18", "resnet50"):
        from . import resnet
        return getattr(resnet, name)(num_classes=num_classes, in_channels=in_channels, small_input=input_hw <= 64)
    raise ValueError(f"unknown model {name!r}; choose from {MODEL_NAMES}")
