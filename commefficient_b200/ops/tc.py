"""Tensor-core layers on the hand-written tcgen05 / TMEM / TMA kernels (``csrc/tc_gemm.cu``,
``csrc/tc_wgrad.cu``): functional entry points plus the autograd Functions the model hooks in
``ops/nn.py`` dispatch to when ``--native_nn`` is on.

Data layout contract: activations are NHWC bf16 (a channels_last ``(N, C, H, W)`` torch tensor IS
an NHWC buffer), convolution weights are ``[Cout][3][3][Cin]`` ("KRSC") -- the layout the flat
parameter buffer uses with ``layout="nhwc"`` -- and linear weights are ``[out][in]``.
Reference layers: /root/reference/CommEfficient/models/resnet9.py:36-37 (conv3x3 s1 p1, no bias).
"""
from __future__ import annotations

import torch

from . import native

native.register_signatures({
    "ceb_gemm_bf16_tn": ([native.c_ptr, native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32,
                          native.c_ptr, native.c_i32, native.c_i32, native.c_ptr], native.c_i32),
    "ceb_conv3x3_bf16_nhwc": ([native.c_ptr, native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32,
                               native.c_i32, native.c_i32, native.c_ptr, native.c_i32, native.c_i32,
                               native.c_ptr], native.c_i32),
    "ceb_conv3x3_wgrad_bf16_nhwc": ([native.c_ptr, native.c_ptr, native.c_ptr, native.c_i32, native.c_i32,
                                     native.c_i32, native.c_i32, native.c_i32, native.c_ptr], native.c_i32),
})

_OUT = {torch.bfloat16: 0, torch.float32: 1}


def gemm_tn(a: torch.Tensor, b: torch.Tensor, bias=None, relu=False, out_dtype=torch.bfloat16, out=None,
            accumulate=False) -> torch.Tensor:
    """``act(a @ b.T + bias)`` with a ``[M, K]``, b ``[N, K]`` contiguous bf16 CUDA tensors."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.is_contiguous() and b.is_contiguous()
    M, K = a.shape
    N, K2 = b.shape
    assert K == K2
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    mode = 2 if accumulate else _OUT[out.dtype]
    native.call("ceb_gemm_bf16_tn", native.ptr(a), native.ptr(b), native.ptr(out), M, N, K, native.ptr(bias),
                1 if relu else 0, mode, native.stream_ptr(a.device))
    return out


def conv3x3_nhwc(x: torch.Tensor, w: torch.Tensor, relu=False, out_dtype=torch.bfloat16) -> torch.Tensor:
    """3x3 / stride 1 / pad 1 convolution.  ``x``: ``[N, H, W, Cin]`` bf16 contiguous, ``w``:
    ``[Cout, 3, 3, Cin]`` bf16 contiguous; returns ``[N, H, W, Cout]``."""
    assert x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and x.is_contiguous() and w.is_contiguous()
    n, h, wd, cin = x.shape
    cout = w.shape[0]
    y = torch.empty((n, h, wd, cout), dtype=out_dtype, device=x.device)
    native.call("ceb_conv3x3_bf16_nhwc", native.ptr(x), native.ptr(w), native.ptr(y), n, h, wd, cin, cout, None,
                1 if relu else 0, _OUT[out_dtype], native.stream_ptr(x.device))
    return y


def conv3x3_wgrad_nhwc(dz: torch.Tensor, x: torch.Tensor, dw: torch.Tensor) -> torch.Tensor:
    """``dw[Cout,3,3,Cin] (fp32, contiguous) += wgrad(dz[N,H,W,Cout], x[N,H,W,Cin])`` (bf16 inputs)."""
    assert dz.dtype == torch.bfloat16 and x.dtype == torch.bfloat16 and dz.is_contiguous() and x.is_contiguous()
    assert dw.dtype == torch.float32 and dw.is_contiguous()
    n, h, wd, cin = x.shape
    cout = dz.shape[3]
    native.call("ceb_conv3x3_wgrad_bf16_nhwc", native.ptr(dz), native.ptr(x), native.ptr(dw), n, h, wd, cin, cout,
                native.stream_ptr(x.device))
    return dw


def conv_supported(x, weight, stride, padding) -> bool:
    return False


def linear_supported(x, weight) -> bool:
    return False
