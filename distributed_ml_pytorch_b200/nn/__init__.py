"""Kernel-backed ``nn.Module`` layers (see :mod:`.layers`)."""
from .layers import (BatchNorm2d, Conv2d, Dropout, Dropout2d, Flatten, GlobalAvgPool, Linear, MaxPool2d, ReLU, ShadowSet, attach_shadows,
                     compute_dtype_name, resolve_dtype, set_compute_dtype)

__all__ = ["Conv2d", "Linear", "ReLU", "MaxPool2d", "Dropout", "Dropout2d", "Flatten", "BatchNorm2d", "GlobalAvgPool", "ShadowSet",
           "attach_shadows", "set_compute_dtype", "compute_dtype_name", "resolve_dtype"]
