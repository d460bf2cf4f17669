"""LeNet (reference example/models.py:5-23), parameter-compatible with the reference ``state_dict``.

The reference computes ``relu(max_pool2d(conv(x), 2))`` and, for the second block,
``relu(max_pool2d(dropout2d(conv(x)), 2))`` (models.py:16-17).  Because ReLU is monotone and Dropout2d scales a
whole channel by 0 or 1/(1-p) >= 0, both equal ``dropout2d(max_pool2d(relu(conv(x))))``, which is what the
fused conv+ReLU+pool block computes here (one GEMM epilogue + one pool kernel instead of three ATen ops).

NORTH-STAR delta: ``in_channels`` / ``input_hw`` make the MNIST shape (1x28x28) legal; the
defaults (3, 32) reproduce the reference exactly.
"""
from __future__ import annotations

import torch.nn as nn

from .. import nn as dnn
from ..ops import functional as OF


class LeNet(nn.Module):
    def __init__(self, in_channels: int = 3, input_hw: int = 32, num_classes: int = 10, dtype="tf32"):
        super().__init__()
        s1 = (input_hw - 4) // 2
        s2 = (s1 - 4) // 2
        self.feat_hw = s2
        self.conv1 = dnn.Conv2d(in_channels, 6, kernel_size=5, activation="relu", pool=(2, 2, 0))
        self.conv2 = dnn.Conv2d(6, 16, kernel_size=5, activation="relu", pool=(2, 2, 0))
        self.conv2_drop = dnn.Dropout2d()
        self.fc1 = dnn.Linear(16 * s2 * s2, 120, activation="relu", in_map=(16, s2, s2))
        self.fc2 = dnn.Linear(120, 84, activation="relu")
        self.fc3 = dnn.Linear(84, num_classes)
        self.fc_drop = dnn.Dropout()  # stateless (F.dropout in the reference, models.py:20)
        dnn.attach_shadows(self)
        # CUDA compute dtype: "tf32" = the reference's precision (fp32 tensors, TF32 tensor-core math), "bf16" = bf16 operands / activations
        dnn.set_compute_dtype(self, dtype)

    def forward(self, x):
        x = self.conv1(x)
        x = self.conv2_drop(self.conv2(x))
        x = OF.flatten(x)
        x = self.fc1(x)
        x = self.fc_drop(x)
        x = self.fc2(x)
        return self.fc3(x)
