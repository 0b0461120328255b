"""Self-contained stand-in for ``fixup.imagenet.models.fixup_resnet_imagenet`` (the external Fixup package the
reference's models/fixup_resnet.py:4 imports and that is not installed here): the bottleneck block and the
ImageNet-style trunk of the Fixup recipe (Zhang et al. 2019), written for this baseline harness only.  Nothing of
commefficient_b200 is imported, so a reference process that builds ``FixupResNet50`` runs reference + shim code
exclusively."""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F


def _conv(cin, cout, k, stride=1):
    return nn.Conv2d(cin, cout, k, stride=stride, padding=k // 2, bias=False)


def _zero():
    return nn.Parameter(torch.zeros(1))


class FixupBottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.bias1a, self.bias1b = _zero(), _zero()
        self.bias2a, self.bias2b = _zero(), _zero()
        self.bias3a, self.bias3b = _zero(), _zero()
        self.conv1 = _conv(inplanes, planes, 1)
        self.conv2 = _conv(planes, planes, 3, stride)
        self.conv3 = _conv(planes, planes * self.expansion, 1)
        self.scale = nn.Parameter(torch.ones(1))
        self.downsample = downsample

    def forward(self, x):
        out = F.relu(self.conv1(x + self.bias1a) + self.bias1b)
        out = F.relu(self.conv2(out + self.bias2a) + self.bias2b)
        out = self.conv3(out + self.bias3a) * self.scale + self.bias3b
        skip = x if self.downsample is None else self.downsample(x + self.bias1a)
        return F.relu(out + skip)


class FixupResNet(nn.Module):
    def __init__(self, block, layers, num_classes=1000):
        super().__init__()
        self.num_layers = sum(layers)
        self.inplanes = 64
        self.conv1 = nn.Conv2d(3, 64, 7, stride=2, padding=3, bias=False)
        self.bias1 = _zero()
        self.maxpool = nn.MaxPool2d(3, stride=2, padding=1)
        self.layer1 = self._stage(block, 64, layers[0], 1)
        self.layer2 = self._stage(block, 128, layers[1], 2)
        self.layer3 = self._stage(block, 256, layers[2], 2)
        self.layer4 = self._stage(block, 512, layers[3], 2)
        self.avgpool = nn.AdaptiveAvgPool2d(1)
        self.bias2 = _zero()
        self.fc = nn.Linear(512 * block.expansion, num_classes)
        for m in self.modules():
            if isinstance(m, FixupBottleneck):
                for conv in (m.conv1, m.conv2):
                    std = math.sqrt(2.0 / (conv.weight.shape[0] * conv.weight.shape[2] * conv.weight.shape[3]))
                    nn.init.normal_(conv.weight, 0.0, std * self.num_layers ** -0.25)
                nn.init.zeros_(m.conv3.weight)
                if m.downsample is not None:
                    w = m.downsample.weight
                    nn.init.normal_(w, 0.0, math.sqrt(2.0 / (w.shape[0] * w.shape[2] * w.shape[3])))
        nn.init.zeros_(self.fc.weight)
        nn.init.zeros_(self.fc.bias)

    def _stage(self, block, planes, n, stride):
        down = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            down = _conv(self.inplanes, planes * block.expansion, 1, stride)
        blocks = [block(self.inplanes, planes, stride, down)]
        self.inplanes = planes * block.expansion
        blocks += [block(self.inplanes, planes) for _ in range(1, n)]
        return nn.Sequential(*blocks)

    def forward(self, x):
        x = self.maxpool(F.relu(self.conv1(x) + self.bias1))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = self.avgpool(x).flatten(1)
        return self.fc(x + self.bias2)
