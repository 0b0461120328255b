"""ImageNet-style residual networks adapted to 1-channel 28x28 FEMNIST input, with
either BatchNorm or LayerNorm (fixed spatial extents 14/7/4/2), plus the ResNeXt /
wide variants.

Capability parity with /root/reference/CommEfficient/models/resnets.py:132-370
(stem ``conv1`` takes one input channel, ``norm_layer=nn.LayerNorm`` supported with
hard-wired feature-map sizes).  ``state_dict`` keys follow the torchvision naming
(``conv1``, ``bn1``, ``layer{1..4}.{i}.conv{1..3}``, ``downsample.0/1``, ``fc``).
Unlike the reference, the two-conv ``BasicBlock`` also accepts the spatial size,
so ``resnet18``/``resnet34`` work (there ``hw`` lands in ``stride``).
No pretrained-weight download: there is no network, and the reference's import of
``torchvision.models.utils`` no longer exists.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Type

import torch
import torch.nn as nn

from ..ops import nn as nnops

__all__ = ["ResNet", "BasicBlock", "Bottleneck", "resnet18", "resnet34", "resnet50", "resnet101",
           "resnet152", "resnext50_32x4d", "resnext101_32x8d", "wide_resnet50_2", "wide_resnet101_2"]


def _norm(norm_layer, channels: int, hw: int):
    """BatchNorm2d(channels) or LayerNorm((channels, hw, hw))."""
    if norm_layer is nn.LayerNorm:
        return nn.LayerNorm((channels, hw, hw))
    return norm_layer(channels)


def _conv(c_in, c_out, k, stride=1, groups=1, dilation=1):
    pad = dilation if k == 3 else 0
    return nn.Conv2d(c_in, c_out, kernel_size=k, stride=stride, padding=pad, groups=groups,
                     dilation=dilation, bias=False)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes, planes, hw, stride=1, downsample=None, groups=1, base_width=64,
                 dilation=1, norm_layer=nnops.BatchNorm2d):
        super().__init__()
        if groups != 1 or base_width != 64 or dilation != 1:
            raise ValueError("BasicBlock supports only groups=1, base_width=64, dilation=1")
        out_hw = math.ceil(hw / stride)
        self.conv1 = _conv(inplanes, planes, 3, stride)
        self.bn1 = _norm(norm_layer, planes, out_hw)
        self.conv2 = _conv(planes, planes, 3)
        self.bn2 = _norm(norm_layer, planes, out_hw)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        skip = x if self.downsample is None else self.downsample(x)
        y = self.relu(self.bn1(self.conv1(x)))
        y = self.bn2(self.conv2(y))
        return self.relu(y + skip)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, hw, stride=1, downsample=None, groups=1, base_width=64,
                 dilation=1, norm_layer=nnops.BatchNorm2d):
        super().__init__()
        width = int(planes * (base_width / 64.0)) * groups
        out_hw = math.ceil(hw / stride)
        self.conv1 = _conv(inplanes, width, 1)
        self.bn1 = _norm(norm_layer, width, hw)
        self.conv2 = _conv(width, width, 3, stride, groups, dilation)
        self.bn2 = _norm(norm_layer, width, out_hw)
        self.conv3 = _conv(width, planes * self.expansion, 1)
        self.bn3 = _norm(norm_layer, planes * self.expansion, out_hw)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        skip = x if self.downsample is None else self.downsample(x)
        y = self.relu(self.bn1(self.conv1(x)))
        y = self.relu(self.bn2(self.conv2(y)))
        y = self.bn3(self.conv3(y))
        return self.relu(y + skip)


class ResNet(nn.Module):
    # feature-map edge entering each stage for a 28x28 input (stem /2, maxpool /2)
    STAGE_HW = (7, 7, 4, 2)
    STAGE_PLANES = (64, 128, 256, 512)

    def __init__(self, block: Type[nn.Module], layers: Sequence[int], num_classes=1000,
                 zero_init_residual=False, groups=1, width_per_group=64,
                 replace_stride_with_dilation: Optional[Sequence[bool]] = None, norm_layer=None,
                 in_channels=1):
        super().__init__()
        norm_layer = norm_layer or nnops.BatchNorm2d
        self._norm_layer = norm_layer
        dilate = list(replace_stride_with_dilation or (False, False, False))
        if len(dilate) != 3:
            raise ValueError("replace_stride_with_dilation needs three entries")
        self.inplanes, self.dilation = 64, 1
        self.groups, self.base_width = groups, width_per_group

        self.conv1 = nn.Conv2d(in_channels, 64, kernel_size=7, stride=2, padding=3, bias=False)
        self.bn1 = _norm(norm_layer, 64, 14)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(kernel_size=3, stride=2, padding=1)
        stages: List[nn.Module] = []
        for i, (planes, n_blocks, hw) in enumerate(zip(self.STAGE_PLANES, layers, self.STAGE_HW)):
            stride = 1 if i == 0 else 2
            stages.append(self._make_stage(block, planes, n_blocks, hw, stride,
                                           dilate[i - 1] if i > 0 else False))
        self.layer1, self.layer2, self.layer3, self.layer4 = stages
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Linear(512 * block.expansion, num_classes)

        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, (nn.BatchNorm2d, nn.GroupNorm)):
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)
        if zero_init_residual:
            for m in self.modules():
                last = getattr(m, "bn3", None) if isinstance(m, Bottleneck) else (
                    getattr(m, "bn2", None) if isinstance(m, BasicBlock) else None)
                if last is not None:
                    nn.init.zeros_(last.weight)

    def _make_stage(self, block, planes, n_blocks, hw, stride, dilate):
        prev_dilation = self.dilation
        if dilate:
            self.dilation *= stride
            stride = 1
        out_ch = planes * block.expansion
        downsample = None
        if stride != 1 or self.inplanes != out_ch:
            downsample = nn.Sequential(_conv(self.inplanes, out_ch, 1, stride),
                                       _norm(self._norm_layer, out_ch, math.ceil(hw / stride)))
        blocks = [block(self.inplanes, planes, hw, stride, downsample, self.groups, self.base_width,
                        prev_dilation, self._norm_layer)]
        self.inplanes = out_ch
        hw = math.ceil(hw / stride)
        for _ in range(1, n_blocks):
            blocks.append(block(self.inplanes, planes, hw, groups=self.groups,
                                base_width=self.base_width, dilation=self.dilation,
                                norm_layer=self._norm_layer))
        return nn.Sequential(*blocks)

    def forward(self, x):
        x = self.maxpool(self.relu(self.bn1(self.conv1(x))))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        return self.fc(torch.flatten(self.avgpool(x), 1))


_SPECS = {
    "resnet18": (BasicBlock, (2, 2, 2, 2), {}),
    "resnet34": (BasicBlock, (3, 4, 6, 3), {}),
    "resnet50": (Bottleneck, (3, 4, 6, 3), {}),
    "resnet101": (Bottleneck, (3, 4, 23, 3), {}),
    "resnet152": (Bottleneck, (3, 8, 36, 3), {}),
    "resnext50_32x4d": (Bottleneck, (3, 4, 6, 3), dict(groups=32, width_per_group=4)),
    "resnext101_32x8d": (Bottleneck, (3, 4, 23, 3), dict(groups=32, width_per_group=8)),
    "wide_resnet50_2": (Bottleneck, (3, 4, 6, 3), dict(width_per_group=128)),
    "wide_resnet101_2": (Bottleneck, (3, 4, 23, 3), dict(width_per_group=128)),
}


def _factory(name):
    block, layers, extra = _SPECS[name]

    def build(pretrained=False, progress=True, **kwargs):
        if pretrained:
            raise RuntimeError("pretrained weights cannot be downloaded in this environment")
        return ResNet(block, layers, **{**extra, **kwargs})

    build.__name__ = name
    build.__doc__ = "%s for 1-channel 28x28 input (see ResNet)." % name
    return build


resnet18 = _factory("resnet18")
resnet34 = _factory("resnet34")
resnet50 = _factory("resnet50")
resnet101 = _factory("resnet101")
resnet152 = _factory("resnet152")
resnext50_32x4d = _factory("resnext50_32x4d")
resnext101_32x8d = _factory("resnext101_32x8d")
wide_resnet50_2 = _factory("wide_resnet50_2")
wide_resnet101_2 = _factory("wide_resnet101_2")
