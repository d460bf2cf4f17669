"""Kernel-backed ``nn.Module`` layers (see :mod:`.layers`)."""
from .layers import (Conv2d, Dropout, Dropout2d, Flatten, Linear, MaxPool2d, ReLU, ShadowSet, attach_shadows)

__all__ = ["Conv2d", "Linear", "ReLU", "MaxPool2d", "Dropout", "Dropout2d", "Flatten", "ShadowSet", "attach_shadows"]
