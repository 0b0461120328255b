"""Functional hooks the models call for their GEMM-shaped layers.

Two implementations sit behind each hook:
  * the library path (``torch.nn.functional`` -> cuDNN / cuBLAS), always available,
    used on CPU and as the numerics reference;
  * the native path (hand-written tcgen05/TMEM/TMA kernels in ``csrc/``, wrapped as
    autograd Functions in ``ops/tc.py``), enabled with ``set_native(True)`` /
    ``--native_nn`` on an sm_100a device.

The hooks keep parameter names/layouts untouched, so switching paths never changes
a checkpoint.  Reference layers: /root/reference/CommEfficient/models/resnet9.py:36-37
(3x3 convs), fixup_resnet18.py:144-154 (3x3 s1/s2, 1x1 s2), GPT-2 linears.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

_native = False


def set_native(flag: bool) -> None:
    global _native
    _native = bool(flag)


def native_enabled() -> bool:
    return _native


def _use_native(x: torch.Tensor) -> bool:
    return _native and x.is_cuda


def conv2d(x, weight, bias=None, stride=1, padding=0):
    if _use_native(x):
        from . import tc
        if tc.conv_supported(x, weight, stride, padding):
            return tc.conv2d(x, weight, bias, stride, padding, relu=False)
    return F.conv2d(x, weight, bias, stride=stride, padding=padding)


def conv2d_relu(x, weight, bias=None, stride=1, padding=0):
    if _use_native(x):
        from . import tc
        if tc.conv_supported(x, weight, stride, padding):
            return tc.conv2d(x, weight, bias, stride, padding, relu=True)
    return F.relu(F.conv2d(x, weight, bias, stride=stride, padding=padding))


def linear(x, weight, bias=None):
    if _use_native(x):
        from . import tc
        if tc.linear_supported(x, weight):
            return tc.linear(x, weight, bias)
    return F.linear(x, weight, bias)
