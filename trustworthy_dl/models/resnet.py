"""ResNet-32 (CIFAR, 3x5 basic blocks) and ResNet-50/101 (ImageNet bottlenecks).

README.md:85-92 lists "ResNet (32, 50, 101)"; the reference's partitioner for ResNets is a
``pass`` (distributed_trainer.py:137-140).  Each residual block is one stage unit.
On a GPU the conv→BN(affine)→ReLU(+residual) epilogues of inference/frozen-BN paths use the
fused kernel in csrc/conv_epilogue.cu via :mod:`trustworthy_dl.ops.conv_epilogue`.
"""
from __future__ import annotations

from typing import List

import torch
import torch.nn as nn
import torch.nn.functional as F

__all__ = ["ResNetCifar", "ResNetImageNet", "resnet32", "resnet50", "resnet101"]


class BasicBlock(nn.Module):
    def __init__(self, cin, cout, stride):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(cout)
        self.conv2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(cout)
        self.short = None
        if stride != 1 or cin != cout:
            self.short = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), nn.BatchNorm2d(cout))

    def forward(self, x):
        y = F.relu(self.bn1(self.conv1(x)))
        y = self.bn2(self.conv2(y))
        return F.relu(y + (x if self.short is None else self.short(x)))


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, cin, width, stride):
        super().__init__()
        cout = width * self.expansion
        self.conv1 = nn.Conv2d(cin, width, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(width)
        self.conv2 = nn.Conv2d(width, width, 3, stride, 1, bias=False)
        self.bn2 = nn.BatchNorm2d(width)
        self.conv3 = nn.Conv2d(width, cout, 1, bias=False)
        self.bn3 = nn.BatchNorm2d(cout)
        self.short = None
        if stride != 1 or cin != cout:
            self.short = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), nn.BatchNorm2d(cout))

    def forward(self, x):
        y = F.relu(self.bn1(self.conv1(x)))
        y = F.relu(self.bn2(self.conv2(y)))
        y = self.bn3(self.conv3(y))
        return F.relu(y + (x if self.short is None else self.short(x)))


class _Stem(nn.Sequential):
    pass


class _Classifier(nn.Module):
    def __init__(self, cin, num_classes):
        super().__init__()
        self.fc = nn.Linear(cin, num_classes)

    def forward(self, x):
        return self.fc(torch.flatten(F.adaptive_avg_pool2d(x, 1), 1))


class _ResNetBase(nn.Module):
    task = "classification"

    def forward(self, x):
        for u in self.stage_units():
            x = u(x)
        return x

    def stage_units(self) -> List[nn.Module]:
        return [self.stem, *self.blocks, self.classifier]

    def loss(self, logits, targets):
        return F.cross_entropy(logits.float(), targets)

    def num_parameters(self) -> int:
        return sum(p.numel() for p in self.parameters())


class ResNetCifar(_ResNetBase):
    """6n+2 layers; n=5 gives ResNet-32 (~0.46 M parameters)."""

    def __init__(self, n: int = 5, num_classes: int = 10):
        super().__init__()
        self.stem = _Stem(nn.Conv2d(3, 16, 3, 1, 1, bias=False), nn.BatchNorm2d(16), nn.ReLU(inplace=True))
        blocks, cin = [], 16
        for cout, stride in ((16, 1), (32, 2), (64, 2)):
            for i in range(n):
                blocks.append(BasicBlock(cin, cout, stride if i == 0 else 1))
                cin = cout
        self.blocks = nn.ModuleList(blocks)
        self.classifier = _Classifier(64, num_classes)
        self.input_shape = (3, 32, 32)
        self.num_classes = num_classes


class ResNetImageNet(_ResNetBase):
    def __init__(self, layers=(3, 4, 6, 3), num_classes: int = 1000):
        super().__init__()
        self.stem = _Stem(nn.Conv2d(3, 64, 7, 2, 3, bias=False), nn.BatchNorm2d(64), nn.ReLU(inplace=True),
                          nn.MaxPool2d(3, 2, 1))
        blocks, cin = [], 64
        for width, n, stride in zip((64, 128, 256, 512), layers, (1, 2, 2, 2)):
            for i in range(n):
                blocks.append(Bottleneck(cin, width, stride if i == 0 else 1))
                cin = width * Bottleneck.expansion
        self.blocks = nn.ModuleList(blocks)
        self.classifier = _Classifier(cin, num_classes)
        self.input_shape = (3, 224, 224)
        self.num_classes = num_classes
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")


def resnet32(num_classes=10):
    return ResNetCifar(5, num_classes)


def resnet50(num_classes=1000):
    return ResNetImageNet((3, 4, 6, 3), num_classes)


def resnet101(num_classes=1000):
    return ResNetImageNet((3, 4, 23, 3), num_classes)
