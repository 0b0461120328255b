"""ResNet-9 (cifar10-fast / DAWNBench lineage).

Architecture and ``state_dict`` key names match
/root/reference/CommEfficient/models/resnet9.py:74-145 so checkpoints are
interchangeable: ``n.prep.conv.weight``, ``n.layer1.conv.weight``,
``n.res1.res1.conv.weight`` ... ``n.linear.weight`` (BatchNorm adds ``.bn.*``).

  prep(c_in->64) . layer1(64->128, pool2) + res1 . layer2(128->256, pool2)
  . layer3(256->512, pool2) + res3 . maxpool4 . linear(512->classes, no bias) . x0.125

All convolutions are 3x3 / stride 1 / pad 1 / no bias; BatchNorm is optional and
off by default (reference default, utils.py:138).  d = 6 568 640 for 10 classes.

The forward is expressed through ``ops.nn`` functional hooks so the hand-written
tcgen05 implicit-GEMM convolution can be swapped in (``set_native``) without
changing parameter names or layout.
"""
from __future__ import annotations

import itertools
from typing import Dict, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import nn as nnops

__all__ = ["ResNet9"]

DEFAULT_CHANNELS = {"prep": 64, "layer1": 128, "layer2": 256, "layer3": 512}


def _make_bn(channels, bn_bias_init=None, bn_bias_freeze=False, bn_weight_init=None,
             bn_weight_freeze=False):
    bn = nnops.BatchNorm2d(channels)
    if bn_bias_init is not None:
        nn.init.constant_(bn.bias, bn_bias_init)
    if bn_weight_init is not None:
        nn.init.constant_(bn.weight, bn_weight_init)
    bn.bias.requires_grad = not bn_bias_freeze
    bn.weight.requires_grad = not bn_weight_freeze
    return bn


class ConvBN(nn.Module):
    """conv3x3 -> [BatchNorm] -> ReLU -> [pool]."""

    def __init__(self, do_batchnorm, c_in, c_out, bn_weight_init=1.0, pool=None, **bn_kw):
        super().__init__()
        self.conv = nn.Conv2d(c_in, c_out, kernel_size=3, stride=1, padding=1, bias=False)
        self.do_batchnorm = bool(do_batchnorm)
        if self.do_batchnorm:
            self.bn = _make_bn(c_out, bn_weight_init=bn_weight_init, **bn_kw)
        self.pool = pool

    def forward(self, x):
        if self.do_batchnorm:
            y = F.relu(self.bn(nnops.conv2d(x, self.conv.weight, stride=1, padding=1)))
        else:
            y = nnops.conv2d_relu(x, self.conv.weight, stride=1, padding=1)
        if self.pool is not None:
            y = self.pool(y)
        return y


class Residual(nn.Module):
    """x + relu(res2(res1(x))) (the inner blocks already end in ReLU)."""

    def __init__(self, do_batchnorm, c, **kw):
        super().__init__()
        self.res1 = ConvBN(do_batchnorm, c, c, **kw)
        self.res2 = ConvBN(do_batchnorm, c, c, **kw)

    def forward(self, x):
        return x + F.relu(self.res2(self.res1(x)))


class BasicNet(nn.Module):
    def __init__(self, do_batchnorm, channels, weight, pool, num_classes, initial_channels=3,
                 new_num_classes=None, **kw):
        super().__init__()
        self.new_num_classes = new_num_classes
        self.channels = channels
        self.prep = ConvBN(do_batchnorm, initial_channels, channels["prep"], **kw)
        self.layer1 = ConvBN(do_batchnorm, channels["prep"], channels["layer1"], pool=pool, **kw)
        self.res1 = Residual(do_batchnorm, channels["layer1"], **kw)
        self.layer2 = ConvBN(do_batchnorm, channels["layer1"], channels["layer2"], pool=pool, **kw)
        self.layer3 = ConvBN(do_batchnorm, channels["layer2"], channels["layer3"], pool=pool, **kw)
        self.res3 = Residual(do_batchnorm, channels["layer3"], **kw)
        self.pool = nn.MaxPool2d(4)
        self.linear = nn.Linear(channels["layer3"], num_classes, bias=False)
        self.out_scale = float(weight)

    def forward(self, x):
        y = self.prep(x)
        y = self.res1(self.layer1(y))
        y = self.layer2(y)
        y = self.res3(self.layer3(y))
        # MaxPool2d(4) over the final 4x4 map of a 32x32 input; for other input sizes (FEMNIST's
        # 28x28 ends at 3x3, where the reference's fixed 4x4 window raises) pool the whole map.
        y = (self.pool(y) if y.shape[-1] == 4 and y.shape[-2] == 4 else F.adaptive_max_pool2d(y, 1)).flatten(1)
        return nnops.linear(y, self.linear.weight, None) * self.out_scale

    def finetune_parameters(self):
        """Fresh classification head; everything else stays frozen (the caller froze
        it).  Reference: resnet9.py:105-113."""
        classes = self.new_num_classes
        if classes is None:
            raise ValueError("finetuning needs new_num_classes")
        self.linear = nn.Linear(self.channels["layer3"], classes, bias=False)
        for p in self.linear.parameters():
            p.requires_grad = True
        return self.linear.parameters()


class ResNet9(nn.Module):
    def __init__(self, do_batchnorm=False, channels: Optional[Dict[str, int]] = None, weight=0.125,
                 pool=None, extra_layers=(), res_layers=("layer1", "layer3"), **kw):
        super().__init__()
        self.channels = dict(channels or DEFAULT_CHANNELS)
        self.weight = weight
        self.pool = pool if pool is not None else nn.MaxPool2d(2)
        self.n = BasicNet(do_batchnorm, self.channels, weight, self.pool, **kw)
        self.kw = kw

    def forward(self, x):
        if nnops.native_enabled():
            from ..ops import tc
            if tc.resnet9_supported(self, x):
                return tc.resnet9_forward(self, x)
        return self.n(x)

    def ce_per_sample(self, x, targets):
        """Per-sample ``(cross-entropy, top-1 correct)``.  On the native path the classifier head and
        the loss are fused with the trunk's autograd Function (``ops/tc.py:_ResNet9CEFn``)."""
        if nnops.native_enabled() and torch.is_grad_enabled():
            from ..ops import tc
            if tc.resnet9_supported(self, x) and tc.head_supported(self.n, self.n.channels["layer3"]):
                return tc.resnet9_ce_per_sample(self, x, targets)
        logits = self(x).float()
        return (F.cross_entropy(logits, targets, reduction="none"), (logits.argmax(dim=1) == targets).float())

    def finetune_parameters(self):
        # (the reference version dereferences an undefined ``self.iid`` -- SURVEY Q-list)
        return self.n.finetune_parameters()
