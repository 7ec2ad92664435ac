"""ResNet-32 (CIFAR, 3x5 basic blocks) and ResNet-50/101 (ImageNet bottlenecks).

README.md:85-92 lists "ResNet (32, 50, 101)"; the reference's partitioner for ResNets is a
``pass`` (distributed_trainer.py:137-140).  Each residual block is one stage unit.
On a GPU every BatchNorm (+ residual add) (+ ReLU) that follows a convolution is ONE pass of the conv-epilogue
kernel (csrc/conv_epilogue.cu, K13, via :mod:`trustworthy_dl.ops.conv_epilogue`), in training as well as in
evaluation: the per-channel batch statistics come from two fp32 reductions over the conv output, are folded into
a scale/shift pair, and the kernel applies scale, shift, residual and activation together (backward: one pass
that yields dx, the residual gradient and the per-channel scale/shift gradients; autograd carries those two back
through the statistics).  Stock eager runs BN, add and ReLU as three kernels (5 reads + 3 writes of the
activation against 3 reads + 1 write here).
"""
from __future__ import annotations

from typing import List

import torch
import torch.nn as nn
import torch.nn.functional as F

__all__ = ["ResNetCifar", "ResNetImageNet", "resnet32", "resnet50", "resnet101"]


class BNAct(nn.BatchNorm2d):
    """``BatchNorm2d`` (same parameters, buffers and state_dict) whose call also takes the residual and the activation:
    ``relu(bn(x) + residual)``.  CUDA tensors go through the fused K13 epilogue, CPU tensors through stock ops."""

    def forward(self, x, residual=None, relu: bool = False):
        if not x.is_cuda or not self.affine or not self.track_running_stats:
            y = super().forward(x)
            if residual is not None:
                y = y + residual
            return F.relu(y) if relu else y
        from ..ops.conv_epilogue import fused_affine_add_relu
        if self.training:
            n = x.numel() // x.shape[1]
            mean = x.mean(dim=(0, 2, 3), dtype=torch.float32)                       # fp32 accumulation, no fp32 copy of x
            ex2 = torch.linalg.vector_norm(x, 2, dim=(0, 2, 3), dtype=torch.float32).square() / n
            var = (ex2 - mean * mean).clamp_min(0.0)
            with torch.no_grad():
                mom = self.momentum if self.momentum is not None else 0.1
                self.running_mean.lerp_(mean.detach().to(self.running_mean.dtype), mom)
                self.running_var.lerp_((var.detach() * (n / max(1, n - 1))).to(self.running_var.dtype), mom)
                self.num_batches_tracked += 1
        else:
            mean, var = self.running_mean.float(), self.running_var.float()
        scale = self.weight.float() * torch.rsqrt(var + self.eps)
        shift = self.bias.float() - mean * scale
        return fused_affine_add_relu(x, scale, shift, residual, relu)


class BasicBlock(nn.Module):
    def __init__(self, cin, cout, stride):
        super().__init__()
        self.conv1 = nn.Conv2d(cin, cout, 3, stride, 1, bias=False)
        self.bn1 = BNAct(cout)
        self.conv2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False)
        self.bn2 = BNAct(cout)
        self.short = None
        if stride != 1 or cin != cout:
            self.short = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), BNAct(cout))

    def forward(self, x):
        y = self.bn1(self.conv1(x), relu=True)
        return self.bn2(self.conv2(y), residual=x if self.short is None else self.short(x), relu=True)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, cin, width, stride):
        super().__init__()
        cout = width * self.expansion
        self.conv1 = nn.Conv2d(cin, width, 1, bias=False)
        self.bn1 = BNAct(width)
        self.conv2 = nn.Conv2d(width, width, 3, stride, 1, bias=False)
        self.bn2 = BNAct(width)
        self.conv3 = nn.Conv2d(width, cout, 1, bias=False)
        self.bn3 = BNAct(cout)
        self.short = None
        if stride != 1 or cin != cout:
            self.short = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), BNAct(cout))

    def forward(self, x):
        y = self.bn1(self.conv1(x), relu=True)
        y = self.bn2(self.conv2(y), relu=True)
        return self.bn3(self.conv3(y), residual=x if self.short is None else self.short(x), relu=True)


class _Stem(nn.Sequential):
    pass


class _StemBN(BNAct):
    """BN + ReLU of the stem as a plain one-argument module (lives inside an ``nn.Sequential``)."""

    def forward(self, x):
        return super().forward(x, None, True)


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
        self.stem = _Stem(nn.Conv2d(3, 16, 3, 1, 1, bias=False), _StemBN(16))
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
        self.stem = _Stem(nn.Conv2d(3, 64, 7, 2, 3, bias=False), _StemBN(64), nn.MaxPool2d(3, 2, 1))
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
