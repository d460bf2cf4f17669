"""ResNet-18 / ResNet-50 (NORTH-STAR configurations, BASELINE.json configs[2,3]; not part of the reference's
example/models.py).  Module names and parameter shapes follow torchvision, so ``state_dict`` keys are
``conv1.weight``, ``bn1.*``, ``layer{1..4}.{i}.conv{1,2,3}.weight``, ``...bn{1,2,3}.*``, ``...downsample.{0,1}.*``,
``fc.{weight,bias}``: ResNet-18 with ``num_classes=10`` has 11,181,642 parameters in the ImageNet stem
variant, ResNet-50 25,557,032.

Every convolution is im2col + tcgen05 GEMM (1x1 / stride-1 convolutions feed the GEMM directly); batch norm, the residual
add and the ReLU of a block run as one fused normalisation pass (``ops.batch_norm(x, residual=..., relu=True)``).
``small_input=True`` selects the usual CIFAR stem (3x3 stride-1 ``conv1``, no max-pool) for 32x32 images.
"""
from __future__ import annotations

import torch.nn as nn

from .. import nn as dnn


def _conv(cin, cout, k, stride=1):
    return dnn.Conv2d(cin, cout, kernel_size=k, stride=stride, padding=k // 2, bias=False)


class _Downsample(nn.Sequential):
    """``downsample.0`` = 1x1 conv, ``downsample.1`` = batch norm (torchvision naming)."""

    def __init__(self, cin, cout, stride):
        super().__init__(_conv(cin, cout, 1, stride), dnn.BatchNorm2d(cout))


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, cin, planes, stride=1):
        super().__init__()
        self.conv1 = _conv(cin, planes, 3, stride)
        self.bn1 = dnn.BatchNorm2d(planes, activation="relu")
        self.conv2 = _conv(planes, planes, 3)
        self.bn2 = dnn.BatchNorm2d(planes, activation="relu")     # relu(bn2(.) + identity), fused
        self.downsample = _Downsample(cin, planes, stride) if (stride != 1 or cin != planes) else None

    def forward(self, x):
        identity = x if self.downsample is None else self.downsample(x)
        out = self.bn1(self.conv1(x))
        return self.bn2(self.conv2(out), residual=identity)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, cin, planes, stride=1):
        super().__init__()
        self.conv1 = _conv(cin, planes, 1)
        self.bn1 = dnn.BatchNorm2d(planes, activation="relu")
        self.conv2 = _conv(planes, planes, 3, stride)
        self.bn2 = dnn.BatchNorm2d(planes, activation="relu")
        self.conv3 = _conv(planes, planes * 4, 1)
        self.bn3 = dnn.BatchNorm2d(planes * 4, activation="relu")
        self.downsample = _Downsample(cin, planes * 4, stride) if (stride != 1 or cin != planes * 4) else None

    def forward(self, x):
        identity = x if self.downsample is None else self.downsample(x)
        out = self.bn1(self.conv1(x))
        out = self.bn2(self.conv2(out))
        return self.bn3(self.conv3(out), residual=identity)


class ResNet(nn.Module):
    def __init__(self, block, layers, num_classes=1000, in_channels=3, small_input=False):
        super().__init__()
        self.small_input = small_input
        if small_input:
            self.conv1 = dnn.Conv2d(in_channels, 64, kernel_size=3, stride=1, padding=1, bias=False)
        else:
            self.conv1 = dnn.Conv2d(in_channels, 64, kernel_size=7, stride=2, padding=3, bias=False)
        self.bn1 = dnn.BatchNorm2d(64, activation="relu")
        self.maxpool = dnn.MaxPool2d(3, 2, 1)
        self.inplanes = 64
        self.layer1 = self._make_layer(block, 64, layers[0], 1)
        self.layer2 = self._make_layer(block, 128, layers[1], 2)
        self.layer3 = self._make_layer(block, 256, layers[2], 2)
        self.layer4 = self._make_layer(block, 512, layers[3], 2)
        self.avgpool = dnn.GlobalAvgPool()
        self.fc = dnn.Linear(512 * block.expansion, num_classes)
        for m in self.modules():   # torchvision's initialisation
            if isinstance(m, dnn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
        dnn.attach_shadows(self)

    def _make_layer(self, block, planes, n, stride):
        blocks = [block(self.inplanes, planes, stride)]
        self.inplanes = planes * block.expansion
        blocks += [block(self.inplanes, planes) for _ in range(1, n)]
        return nn.Sequential(*blocks)

    def forward(self, x):
        x = self.bn1(self.conv1(x))
        if not self.small_input:
            x = self.maxpool(x)
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        return self.fc(self.avgpool(x))


def resnet18(num_classes=10, in_channels=3, small_input=True):
    return ResNet(BasicBlock, [2, 2, 2, 2], num_classes, in_channels, small_input)


def resnet50(num_classes=1000, in_channels=3, small_input=False):
    return ResNet(Bottleneck, [3, 4, 6, 3], num_classes, in_channels, small_input)
