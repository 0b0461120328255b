"""CIFAR-style 18-layer residual nets: ``ResNet18`` (BatchNorm blocks) and
``FixupResNet18`` (normalisation-free, scalar bias/scale parameters).

Layer widths, strides, head and ``state_dict`` keys follow
/root/reference/CommEfficient/models/fixup_resnet18.py:24-216:
prep 3->64; four stages of two blocks with widths 64/128/256/**256** and strides
1/2/2/2; head = concat(global-avg, global-max) -> Linear(512 -> classes).
d = 5 206 218 (ResNet18, 10 classes) / 5 200 626 (FixupResNet18, 10 classes).
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import nn as nnops

__all__ = ["ResNet18", "FixupResNet18"]

STAGES = ((64, 64, 1), (64, 128, 2), (128, 256, 2), (256, 256, 2))


class Mul(nn.Module):
    def __init__(self):
        super().__init__()
        self.scale = nn.Parameter(torch.ones(1))

    def forward(self, x):
        return x * self.scale


class Add(nn.Module):
    def __init__(self):
        super().__init__()
        self.bias = nn.Parameter(torch.zeros(1))

    def forward(self, x):
        return x + self.bias


def _conv3x3(c_in, c_out, stride):
    return nn.Conv2d(c_in, c_out, kernel_size=3, stride=stride, padding=1, bias=False)


def _needs_projection(c_in, c_out, stride):
    return stride != 1 or c_in != c_out


class FixupBlock(nn.Module):
    """relu( add2b(mul(conv2(add2a( relu(add1b(conv1(add1a(x)))) )))) + shortcut(x) )"""

    def __init__(self, in_channels, out_channels, stride=1):
        super().__init__()
        self.add1a = Add()
        self.conv1 = _conv3x3(in_channels, out_channels, stride)
        self.add1b = Add()
        self.add2a = Add()
        self.conv2 = _conv3x3(out_channels, out_channels, 1)
        self.mul = Mul()
        self.add2b = Add()
        if _needs_projection(in_channels, out_channels, stride):
            self.shortcut = nn.Conv2d(in_channels, out_channels, kernel_size=1, stride=stride, bias=False)

    def forward(self, x):
        skip = x
        if hasattr(self, "shortcut"):
            skip = nnops.conv2d(x, self.shortcut.weight, stride=self.shortcut.stride[0], padding=0)
        y = nnops.conv2d(self.add1a(x), self.conv1.weight, stride=self.conv1.stride[0], padding=1)
        y = F.relu(self.add1b(y))
        y = nnops.conv2d(self.add2a(y), self.conv2.weight, stride=1, padding=1)
        y = self.add2b(self.mul(y))
        return F.relu(y + skip)


class PreActBlock(nn.Module):
    """relu(bn2(conv2(relu(bn1(conv1(x)))))) + shortcut(x); no ReLU after the add and
    no BatchNorm on the 1x1 shortcut (fixup_resnet18.py:138-165)."""

    def __init__(self, in_channels, out_channels, stride=1):
        super().__init__()
        self.bn1 = nnops.BatchNorm2d(out_channels)
        self.conv1 = _conv3x3(in_channels, out_channels, stride)
        self.bn2 = nnops.BatchNorm2d(out_channels)
        self.conv2 = _conv3x3(out_channels, out_channels, 1)
        if _needs_projection(in_channels, out_channels, stride):
            self.shortcut = nn.Sequential(
                nn.Conv2d(in_channels, out_channels, kernel_size=1, stride=stride, bias=False))

    def forward(self, x):
        y = nnops.conv2d(x, self.conv1.weight, stride=self.conv1.stride[0], padding=1)
        y = F.relu(self.bn1(y))
        y = nnops.conv2d(y, self.conv2.weight, stride=1, padding=1)
        y = F.relu(self.bn2(y))
        skip = x
        if hasattr(self, "shortcut"):
            conv = self.shortcut[0]
            skip = nnops.conv2d(x, conv.weight, stride=conv.stride[0], padding=0)
        return y + skip


def _stage(block, c_in, c_out, n_blocks, stride):
    blocks = []
    for i in range(n_blocks):
        blocks.append(block(c_in if i == 0 else c_out, c_out, stride if i == 0 else 1))
    return nn.Sequential(*blocks)


def _avgmax_head(x):
    return torch.cat([F.adaptive_avg_pool2d(x, 1).flatten(1), F.adaptive_max_pool2d(x, 1).flatten(1)], dim=-1)


class ResNet18(nn.Module):
    def __init__(self, num_blocks=(2, 2, 2, 2), num_classes=10, initial_channels=3, **_unused):
        super().__init__()
        self.prep = nn.Sequential(
            nn.Conv2d(initial_channels, 64, kernel_size=3, stride=1, padding=1, bias=False), nn.ReLU())
        self.layers = nn.Sequential(*[_stage(PreActBlock, ci, co, nb, s)
                                      for (ci, co, s), nb in zip(STAGES, num_blocks)])
        self.classifier = nn.Linear(512, num_classes)

    def forward(self, x):
        x = nnops.conv2d_relu(x, self.prep[0].weight, stride=1, padding=1)
        x = self.layers(x)
        return nnops.linear(_avgmax_head(x), self.classifier.weight, self.classifier.bias)


class FixupResNet18(nn.Module):
    def __init__(self, num_blocks=(2, 2, 2, 2), num_classes=10, initial_channels=3, **_unused):
        super().__init__()
        self.num_layers = sum(num_blocks)
        self.prep = nn.Conv2d(initial_channels, 64, kernel_size=3, stride=1, padding=1, bias=False)
        self.layers = nn.Sequential(*[_stage(FixupBlock, ci, co, nb, s)
                                      for (ci, co, s), nb in zip(STAGES, num_blocks)])
        self.classifier = nn.Linear(512, num_classes)
        self._fixup_init()

    def _fixup_init(self):
        """Fixup initialisation (fixup_resnet18.py:86-106): first conv of each block
        He-normal scaled by L^-1/2, second conv zero, projection He-normal, classifier zero."""
        def he_std(w):
            return math.sqrt(2.0 / (w.shape[0] * w.shape[2] * w.shape[3]))

        for m in self.modules():
            if isinstance(m, FixupBlock):
                nn.init.normal_(m.conv1.weight, 0.0, he_std(m.conv1.weight) * self.num_layers ** -0.5)
                nn.init.zeros_(m.conv2.weight)
                if hasattr(m, "shortcut"):
                    nn.init.normal_(m.shortcut.weight, 0.0, he_std(m.shortcut.weight))
            elif isinstance(m, nn.Linear):
                nn.init.zeros_(m.weight)
                nn.init.zeros_(m.bias)
        nn.init.normal_(self.prep.weight, 0.0, he_std(self.prep.weight))

    def forward(self, x):
        x = nnops.conv2d_relu(x, self.prep.weight, stride=1, padding=1)
        x = self.layers(x)
        return nnops.linear(_avgmax_head(x), self.classifier.weight, self.classifier.bias)
