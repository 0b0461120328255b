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


def layer_norm(x, weight, bias, eps=1e-5):
    if _use_native(x) and weight.is_cuda:
        from . import tfm
        return tfm.layer_norm(x, weight, bias, eps)
    return F.layer_norm(x, (x.shape[-1],), weight, bias, eps)


def gelu_tanh(x):
    if _use_native(x) and x.numel() % 2 == 0:
        from . import tfm
        return tfm.gelu_tanh(x)
    return F.gelu(x, approximate="tanh")


def cross_entropy_rows(logits, labels, ignore_index=-1):
    """Per-row cross-entropy (0 where ``labels == ignore_index``) of ``[rows, V]`` logits."""
    if _use_native(logits) and logits.dtype == torch.bfloat16 and logits.dim() == 2 and logits.stride(1) == 1 \
            and logits.stride(0) % 2 == 0:
        from . import tfm
        return tfm.cross_entropy_rows(logits, labels, ignore_index)
    return F.cross_entropy(logits.float(), labels, ignore_index=ignore_index, reduction="none")


def lm_head_cross_entropy(hidden2d, weight, labels, ignore_index=-1):
    """Per-row cross-entropy of ``hidden2d @ weight.T`` (0 where ``labels == ignore_index``).  Native
    path: one fused function that only computes the rows with a label (``ops/tc.py:_LMHeadCE``)."""
    if _use_native(hidden2d):
        from . import tc
        if tc.lm_head_supported(hidden2d, weight):
            return tc.lm_head_ce(hidden2d, weight, labels, ignore_index)
    return cross_entropy_rows(linear(hidden2d, weight, None), labels, ignore_index)
