"""CIFAR-sized AlexNet (reference example/models.py:25-49), ``state_dict``-compatible with the reference:
``features.{0,3,6,8,10}.{weight,bias}`` and ``classifier.{weight,bias}``.

Every conv runs as im2col + tcgen05 GEMM with bias+ReLU in the epilogue, followed by the 2x2 max-pool
kernel where the reference has one; the ``ReLU`` / ``MaxPool2d`` entries of ``features`` stay in place as
fused placeholders so the ``nn.Sequential`` indices (and therefore checkpoint keys) are unchanged.
"""
from __future__ import annotations

import torch.nn as nn

from .. import nn as dnn
from ..ops import functional as OF


class AlexNet(nn.Module):
    def __init__(self, num_classes: int = 10, in_channels: int = 3, dtype="tf32"):
        super().__init__()
        P = (2, 2, 0)
        self.features = nn.Sequential(
            dnn.Conv2d(in_channels, 64, kernel_size=11, stride=4, padding=5, activation="relu", pool=P, sole_consumer_is_conv=True),
            dnn.ReLU(fused=True),
            dnn.MaxPool2d(2, 2, fused=True),
            dnn.Conv2d(64, 192, kernel_size=5, padding=2, activation="relu", pool=P, sole_consumer_is_conv=True),
            dnn.ReLU(fused=True),
            dnn.MaxPool2d(2, 2, fused=True),
            dnn.Conv2d(192, 384, kernel_size=3, padding=1, activation="relu", sole_consumer_is_conv=True),
            dnn.ReLU(fused=True),
            dnn.Conv2d(384, 256, kernel_size=3, padding=1, activation="relu", sole_consumer_is_conv=True),
            dnn.ReLU(fused=True),
            dnn.Conv2d(256, 256, kernel_size=3, padding=1, activation="relu", pool=P),
            dnn.ReLU(fused=True),
            dnn.MaxPool2d(2, 2, fused=True),
        )
        self.classifier = dnn.Linear(256, num_classes)
        dnn.attach_shadows(self)
        # CUDA compute dtype: "tf32" = the reference's precision (fp32 tensors, TF32 tensor-core math), "bf16" = bf16 operands / activations
        dnn.set_compute_dtype(self, dtype)

    def forward(self, x):
        x = self.features(x)
        x = OF.flatten(x)
        return self.classifier(x)
