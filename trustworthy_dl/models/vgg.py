"""VGG-11/13/16 (README.md:85-92).  Units are conv groups (conv+ReLU runs ending in a max-pool)
and the three classifier layers, so VGG-16 splits naturally into the 8 stages of BASELINE
config 2.  ``input_size`` 32 gives the CIFAR variant (512-d classifier)."""
from __future__ import annotations

from typing import List

import torch
import torch.nn as nn
import torch.nn.functional as F

__all__ = ["VGG", "vgg11", "vgg13", "vgg16"]

_CFG = {
    11: [[64], [128], [256, 256], [512, 512], [512, 512]],
    13: [[64, 64], [128, 128], [256, 256], [512, 512], [512, 512]],
    16: [[64, 64], [128, 128], [256, 256, 256], [512, 512, 512], [512, 512, 512]],
}


class _ConvBiasReLU(nn.Conv2d):
    """3x3 conv whose bias + ReLU run as one fused epilogue kernel on CUDA (csrc/conv_epilogue.cu, K13): the
    convolution itself stays in cuDNN without bias, the epilogue is a single pass over the activation."""

    def forward(self, x):
        if x.is_cuda and self.bias is not None:
            from ..ops.conv_epilogue import fused_bias_relu
            y = F.conv2d(x, self.weight, None, self.stride, self.padding, self.dilation, self.groups)
            return fused_bias_relu(y, self.bias)
        return F.relu(super().forward(x))


class _ConvGroup(nn.Sequential):
    def __init__(self, cin, widths, batch_norm):
        layers = []
        for w in widths:
            if batch_norm:
                layers += [nn.Conv2d(cin, w, 3, padding=1), nn.BatchNorm2d(w), nn.ReLU(inplace=True)]
            else:
                layers.append(_ConvBiasReLU(cin, w, 3, padding=1))
            cin = w
        layers.append(nn.MaxPool2d(2, 2))
        super().__init__(*layers)


class _FC(nn.Module):
    def __init__(self, cin, cout, act=True, flatten=False, drop=0.0):
        super().__init__()
        self.fc = nn.Linear(cin, cout)
        self.act, self.flatten, self.drop = act, flatten, drop

    def forward(self, x):
        if self.flatten:
            x = torch.flatten(x, 1)
        x = self.fc(x)
        if self.act:
            x = F.relu(x)
            if self.drop > 0:
                x = F.dropout(x, self.drop, self.training)
        return x


class VGG(nn.Module):
    task = "classification"

    def __init__(self, depth: int = 16, num_classes: int = 1000, input_size: int = 224,
                 batch_norm: bool = False, dropout: float = 0.0):
        super().__init__()
        groups, cin = [], 3
        for widths in _CFG[depth]:
            groups.append(_ConvGroup(cin, widths, batch_norm))
            cin = widths[-1]
        self.features = nn.ModuleList(groups)
        spatial = input_size // 32
        hidden = 4096 if input_size >= 64 else 512
        self.classifier = nn.ModuleList([
            _FC(512 * spatial * spatial, hidden, flatten=True, drop=dropout),
            _FC(hidden, hidden, drop=dropout),
            _FC(hidden, num_classes, act=False),
        ])
        self.input_shape = (3, input_size, input_size)
        self.num_classes = num_classes
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
                nn.init.zeros_(m.bias)
            elif isinstance(m, nn.Linear):
                nn.init.normal_(m.weight, 0, 0.01)
                nn.init.zeros_(m.bias)

    def stage_units(self) -> List[nn.Module]:
        return [*self.features, *self.classifier]

    def forward(self, x):
        for u in self.stage_units():
            x = u(x)
        return x

    def loss(self, logits, targets):
        return F.cross_entropy(logits.float(), targets)

    def num_parameters(self) -> int:
        return sum(p.numel() for p in self.parameters())


def vgg11(**kw):
    return VGG(11, **kw)


def vgg13(**kw):
    return VGG(13, **kw)


def vgg16(**kw):
    return VGG(16, **kw)
