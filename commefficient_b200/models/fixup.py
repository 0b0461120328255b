"""Fixup (normalisation-free) residual networks: ``FixupResNet9`` and
``FixupResNet50``.

The reference gets these from the external ``fixup`` package, which is not
installed here (/root/reference/CommEfficient/models/fixup_resnet9.py:6,
models/fixup_resnet.py:4); the blocks are therefore defined in this file from the
Fixup recipe (Zhang et al. 2019): scalar biases before every conv/activation, one
scalar multiplier per residual branch, first convs of a branch scaled by
L^(-1/(2m-2)) for an m-conv branch, last conv of a branch zero, classifier zero.
Parameter names contain ``bias``/``scale`` so the driver's Fixup learning-rate
groups (cv_train.py:366-376) pick them up.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

__all__ = ["FixupResNet9", "FixupResNet50", "FixupBasicBlock", "FixupBottleneck", "FixupLayer"]


def _scalar(value=0.0):
    return nn.Parameter(torch.full((1,), float(value)))


def _conv3x3(c_in, c_out, stride=1):
    return nn.Conv2d(c_in, c_out, 3, stride=stride, padding=1, bias=False)


def _conv1x1(c_in, c_out, stride=1):
    return nn.Conv2d(c_in, c_out, 1, stride=stride, bias=False)


def _he_std(w):
    return math.sqrt(2.0 / (w.shape[0] * w.shape[2] * w.shape[3]))


class FixupBasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.bias1a = _scalar()
        self.conv1 = _conv3x3(inplanes, planes, stride)
        self.bias1b = _scalar()
        self.bias2a = _scalar()
        self.conv2 = _conv3x3(planes, planes)
        self.scale = _scalar(1.0)
        self.bias2b = _scalar()
        self.downsample = downsample

    def forward(self, x):
        y = F.relu(self.conv1(x + self.bias1a) + self.bias1b)
        y = self.conv2(y + self.bias2a) * self.scale + self.bias2b
        skip = x if self.downsample is None else self.downsample(x + self.bias1a)
        return F.relu(y + skip)


class FixupBottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.bias1a = _scalar()
        self.conv1 = _conv1x1(inplanes, planes)
        self.bias1b = _scalar()
        self.bias2a = _scalar()
        self.conv2 = _conv3x3(planes, planes, stride)
        self.bias2b = _scalar()
        self.bias3a = _scalar()
        self.conv3 = _conv1x1(planes, planes * self.expansion)
        self.scale = _scalar(1.0)
        self.bias3b = _scalar()
        self.downsample = downsample

    def forward(self, x):
        y = F.relu(self.conv1(x + self.bias1a) + self.bias1b)
        y = F.relu(self.conv2(y + self.bias2a) + self.bias2b)
        y = self.conv3(y + self.bias3a) * self.scale + self.bias3b
        skip = x if self.downsample is None else self.downsample(x + self.bias1a)
        return F.relu(y + skip)


class FixupLayer(nn.Module):
    """conv -> scale/bias -> relu -> [pool] -> ``num_blocks`` FixupBasicBlocks."""

    def __init__(self, in_channels, out_channels, num_blocks, pool):
        super().__init__()
        self.conv = _conv3x3(in_channels, out_channels)
        self.bias1a = _scalar()
        self.bias1b = _scalar()
        self.scale = _scalar(1.0)
        self.pool = pool
        self.blocks = nn.Sequential(*[FixupBasicBlock(out_channels, out_channels) for _ in range(num_blocks)])

    def forward(self, x):
        y = F.relu(self.conv(x + self.bias1a) * self.scale + self.bias1b)
        if self.pool is not None:
            y = self.pool(y)
        return self.blocks(y)


class FixupResNet9(nn.Module):
    """ResNet-9 layout with Fixup blocks (fixup_resnet9.py:33-91).  The reference
    hard-codes 10 classes; ``num_classes`` is a parameter here."""

    def __init__(self, channels=None, pool=None, num_classes=10, initial_channels=3, **_unused):
        super().__init__()
        pool = pool if pool is not None else nn.MaxPool2d(2)
        self.num_layers = 2
        ch = dict(channels or {"prep": 64, "layer1": 128, "layer2": 256, "layer3": 512})
        self.channels = ch
        self.conv1 = _conv3x3(initial_channels, ch["prep"])
        self.bias1a, self.bias1b, self.scale = _scalar(), _scalar(), _scalar(1.0)
        self.layer1 = FixupLayer(ch["prep"], ch["layer1"], 1, pool)
        self.layer2 = FixupLayer(ch["layer1"], ch["layer2"], 0, pool)
        self.layer3 = FixupLayer(ch["layer2"], ch["layer3"], 1, pool)
        self.pool = nn.MaxPool2d(4)
        self.bias2 = _scalar()
        self.linear = nn.Linear(ch["layer3"], num_classes)
        nn.init.normal_(self.conv1.weight, 0.0, _he_std(self.conv1.weight))
        for m in self.modules():
            if isinstance(m, FixupBasicBlock):
                nn.init.normal_(m.conv1.weight, 0.0, _he_std(m.conv1.weight) * self.num_layers ** -0.5)
                nn.init.zeros_(m.conv2.weight)
            elif isinstance(m, FixupLayer):
                nn.init.normal_(m.conv.weight, 0.0, _he_std(m.conv.weight))
            elif isinstance(m, nn.Linear):
                nn.init.zeros_(m.weight)
                nn.init.zeros_(m.bias)

    def forward(self, x):
        y = F.relu(self.conv1(x + self.bias1a) * self.scale + self.bias1b)
        y = self.layer3(self.layer2(self.layer1(y)))
        y = self.pool(y).flatten(1)
        return self.linear(y + self.bias2)


class FixupResNet(nn.Module):
    """ImageNet-layout Fixup ResNet (7x7/2 stem, maxpool, four stages, avgpool, fc)."""

    def __init__(self, block, layers, num_classes=1000, in_channels=3, **_unused):
        super().__init__()
        self.num_layers = sum(layers)
        self.inplanes = 64
        self.conv1 = nn.Conv2d(in_channels, 64, kernel_size=7, stride=2, padding=3, bias=False)
        self.bias1 = _scalar()
        self.maxpool = nn.MaxPool2d(kernel_size=3, stride=2, padding=1)
        self.layer1 = self._stage(block, 64, layers[0], 1)
        self.layer2 = self._stage(block, 128, layers[1], 2)
        self.layer3 = self._stage(block, 256, layers[2], 2)
        self.layer4 = self._stage(block, 512, layers[3], 2)
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.bias2 = _scalar()
        self.fc = nn.Linear(512 * block.expansion, num_classes)
        depth_scale = {FixupBasicBlock: -0.5, FixupBottleneck: -0.25}[block]
        for m in self.modules():
            if isinstance(m, FixupBasicBlock):
                nn.init.normal_(m.conv1.weight, 0.0, _he_std(m.conv1.weight) * self.num_layers ** depth_scale)
                nn.init.zeros_(m.conv2.weight)
            elif isinstance(m, FixupBottleneck):
                nn.init.normal_(m.conv1.weight, 0.0, _he_std(m.conv1.weight) * self.num_layers ** depth_scale)
                nn.init.normal_(m.conv2.weight, 0.0, _he_std(m.conv2.weight) * self.num_layers ** depth_scale)
                nn.init.zeros_(m.conv3.weight)
            elif isinstance(m, nn.Linear):
                nn.init.zeros_(m.weight)
                nn.init.zeros_(m.bias)
        for m in self.modules():
            ds = getattr(m, "downsample", None)
            if isinstance(ds, nn.Conv2d):
                nn.init.normal_(ds.weight, 0.0, _he_std(ds.weight))

    def _stage(self, block, planes, n_blocks, stride):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = _conv1x1(self.inplanes, planes * block.expansion, stride)
        blocks = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        blocks += [block(self.inplanes, planes) for _ in range(1, n_blocks)]
        return nn.Sequential(*blocks)

    def forward(self, x):
        x = self.maxpool(F.relu(self.conv1(x) + self.bias1))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = torch.flatten(self.avgpool(x), 1)
        return self.fc(x + self.bias2)


class FixupResNet50(FixupResNet):
    """fixup_resnet.py:8-10."""

    def __init__(self, **kwargs):
        kwargs.pop("do_batchnorm", None)
        kwargs.pop("channels", None)
        kwargs.pop("new_num_classes", None)
        kwargs.pop("bn_bias_freeze", None)
        kwargs.pop("bn_weight_freeze", None)
        super().__init__(FixupBottleneck, [3, 4, 6, 3], **kwargs)
