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
    "ceb_gemm_bf16_nn": ([native.c_ptr, native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32,
                          native.c_ptr, native.c_i32, native.c_i32, native.c_ptr], native.c_i32),
    "ceb_conv3x3_bf16_nhwc": ([native.c_ptr, native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32,
                               native.c_i32, native.c_i32, native.c_ptr, native.c_i32, native.c_i32,
                               native.c_ptr, native.c_ptr], native.c_i32),
    "ceb_conv3x3_bf16_nhwc_res": ([native.c_ptr, native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32,
                                   native.c_i32, native.c_i32, native.c_ptr, native.c_i32, native.c_ptr, native.c_ptr,
                                   native.c_ptr, native.c_ptr], native.c_i32),
    "ceb_conv_w_dgrad_layout": ([native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_ptr], native.c_i32),
    "ceb_cast_f32_bf16": ([native.c_ptr, native.c_ptr, native.c_i64, native.c_ptr], native.c_i32),
    "ceb_maxpool_nhwc_bf16": ([native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32, native.c_i32,
                               native.c_i32, native.c_ptr], native.c_i32),
    "ceb_maxpool_bwd_nhwc_bf16": ([native.c_ptr, native.c_ptr, native.c_ptr, native.c_i32, native.c_i32,
                                   native.c_i32, native.c_i32, native.c_i32, native.c_i32, native.c_ptr],
                                  native.c_i32),
    "ceb_relu_bwd_bf16": ([native.c_ptr, native.c_ptr, native.c_ptr, native.c_i64, native.c_ptr], native.c_i32),
    "ceb_im2col3x3_c3": ([native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32, native.c_ptr],
                         native.c_i32),
    "ceb_gemm_bf16_mn_acc": ([native.c_ptr, native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32,
                              native.c_i32, native.c_ptr], native.c_i32),
    "ceb_gemm_bf16_tn_ld": ([native.c_ptr, native.c_i64, native.c_ptr, native.c_i64, native.c_ptr, native.c_i64,
                             native.c_i32, native.c_i32, native.c_i32, native.c_ptr, native.c_i32, native.c_i32,
                             native.c_ptr], native.c_i32),
    "ceb_gemm_bf16_tn_gelu": ([native.c_ptr, native.c_i64, native.c_ptr, native.c_i64, native.c_ptr, native.c_ptr,
                               native.c_ptr, native.c_i64, native.c_i32, native.c_i32, native.c_i32, native.c_ptr,
                               native.c_i32, native.c_ptr], native.c_i32),
    "ceb_gemm_bf16_mn_acc_ld": ([native.c_ptr, native.c_i64, native.c_ptr, native.c_i64, native.c_ptr, native.c_i32,
                                 native.c_i32, native.c_i32, native.c_i32, native.c_ptr], native.c_i32),
    "ceb_cast_transpose_f32_bf16": ([native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i64,
                                     native.c_ptr], native.c_i32),
    "ceb_colsum_bf16": ([native.c_ptr, native.c_i64, native.c_i64, native.c_i32, native.c_ptr, native.c_ptr],
                        native.c_i32),
    "ceb_conv_weights_prepare": ([native.c_i32, native.c_ptr, native.c_ptr, native.c_ptr, native.c_ptr, native.c_ptr,
                                  native.c_ptr], native.c_i32),
    "ceb_gemm_bf16_tn_rows": ([native.c_ptr, native.c_i64, native.c_ptr, native.c_i64, native.c_ptr, native.c_i64,
                               native.c_i32, native.c_i32, native.c_i32, native.c_ptr, native.c_i32, native.c_i32,
                               native.c_ptr, native.c_ptr], native.c_i32),
    "ceb_gemm_bf16_mn_acc_rows": ([native.c_ptr, native.c_i64, native.c_ptr, native.c_i64, native.c_ptr, native.c_i32,
                                   native.c_i32, native.c_i32, native.c_i32, native.c_ptr, native.c_ptr],
                                  native.c_i32),
    "ceb_softmax_ce_bwd_rows": ([native.c_ptr, native.c_ptr, native.c_i64, native.c_i64, native.c_i32, native.c_ptr,
                                 native.c_i64, native.c_ptr, native.c_ptr, native.c_ptr, native.c_ptr], native.c_i32),
})

native.register_signatures({
    "ceb_gemm_tf32_tn_ld": ([native.c_ptr, native.c_i64, native.c_ptr, native.c_i64, native.c_ptr, native.c_i64,
                             native.c_i32, native.c_i32, native.c_i32, native.c_ptr, native.c_i32, native.c_i32,
                             native.c_ptr], native.c_i32),
    "ceb_conv3x3_tf32_nhwc": ([native.c_ptr, native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32,
                               native.c_i32, native.c_i32, native.c_ptr, native.c_i32, native.c_ptr, native.c_ptr,
                               native.c_ptr, native.c_ptr], native.c_i32),
    "ceb_conv3x3_wgrad_tf32_nhwc": ([native.c_ptr, native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32,
                                     native.c_i32, native.c_i32, native.c_ptr], native.c_i32),
    "ceb_gemm_tf32_mn_acc_ld": ([native.c_ptr, native.c_i64, native.c_ptr, native.c_i64, native.c_ptr, native.c_i32,
                                 native.c_i32, native.c_i32, native.c_i32, native.c_ptr], native.c_i32),
    "ceb_conv_weights_prepare_f32": ([native.c_i32, native.c_ptr, native.c_ptr, native.c_ptr, native.c_ptr,
                                      native.c_ptr, native.c_ptr], native.c_i32),
    "ceb_maxpool_nhwc_f32": ([native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32, native.c_i32,
                              native.c_i32, native.c_ptr], native.c_i32),
    "ceb_maxpool_bwd_nhwc_f32": ([native.c_ptr, native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32,
                                  native.c_i32, native.c_i32, native.c_i32, native.c_ptr], native.c_i32),
    "ceb_relu_bwd_f32": ([native.c_ptr, native.c_ptr, native.c_ptr, native.c_i64, native.c_ptr], native.c_i32),
    "ceb_im2col3x3": ([native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32, native.c_i32,
                       native.c_i32, native.c_ptr], native.c_i32),
    "ceb_linear_ce_head_f32": ([native.c_ptr, native.c_ptr, native.c_f32, native.c_ptr, native.c_i32, native.c_i32,
                                native.c_i32, native.c_ptr, native.c_ptr, native.c_ptr, native.c_ptr, native.c_ptr,
                                native.c_ptr], native.c_i32),
})

native.register_signatures({
    "ceb_conv2d_nhwc": ([native.c_ptr, native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32, native.c_i32,
                         native.c_i32, native.c_i32, native.c_i32, native.c_i32, native.c_i32, native.c_i32, native.c_ptr,
                         native.c_ptr, native.c_ptr, native.c_ptr], native.c_i32),
    "ceb_conv3x3_wgrad_nhwc": ([native.c_ptr, native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32,
                                native.c_i32, native.c_i32, native.c_i32, native.c_i32, native.c_ptr], native.c_i32),
    "ceb_subsample2_nhwc": ([native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32, native.c_i32,
                             native.c_ptr], native.c_i32),
    "ceb_upsample2_zero_nhwc": ([native.c_ptr, native.c_ptr, native.c_i32, native.c_i32, native.c_i32, native.c_i32,
                                 native.c_i32, native.c_i32, native.c_ptr], native.c_i32),
})

_OUT = {torch.bfloat16: 0, torch.float32: 1}

# Operand precision of the tensor-core layers (``--precision``): "bf16" -> bf16 activations and
# tcgen05.mma.kind::f16; "tf32" -> fp32 activations read by the tensor cores as TF32 (kind::tf32), which is
# what the reference's fp32 convolutions do inside cuDNN (torch.backends.cudnn.allow_tf32 defaults to True).
# "fp32" never reaches this module: ops/nn.py keeps those runs on the library with TF32 switched off.
_PRECISION = "bf16"


def set_precision(p: str) -> None:
    global _PRECISION
    if p not in ("bf16", "tf32"):
        raise ValueError("tensor-core layers run bf16 or tf32 operands, not %r" % (p,))
    _PRECISION = p


def act_dtype() -> torch.dtype:
    return torch.float32 if _PRECISION == "tf32" else torch.bfloat16


def _al(dtype) -> int:
    """Elements per 16 bytes."""
    return 4 if dtype == torch.float32 else 8


def _rows_ok(t: torch.Tensor) -> bool:
    """2-D bf16 / fp32, unit column stride, 16-byte aligned rows."""
    return (t.dim() == 2 and t.dtype in (torch.bfloat16, torch.float32) and t.stride(1) == 1
            and t.stride(0) % _al(t.dtype) == 0 and t.stride(0) >= t.shape[1] and t.data_ptr() % 16 == 0)


def padded_empty(rows: int, cols: int, dtype, device) -> torch.Tensor:
    """``[rows, cols]`` view of a buffer whose row stride is ``cols`` rounded up to 16 elements, so bf16
    rows stay 32-byte aligned (TMA needs 16; the GEMM epilogue stores whole 32-byte sectors) whatever
    ``cols`` is (e.g. the 50 262-wide logits)."""
    ld = (cols + 15) // 16 * 16
    return torch.empty((rows, ld), dtype=dtype, device=device)[:, :cols]


def gemm_tn(a: torch.Tensor, b: torch.Tensor, bias=None, relu=False, out_dtype=torch.bfloat16, out=None,
            accumulate=False) -> torch.Tensor:
    """``act(a @ b.T + bias)`` with a ``[M, K]``, b ``[N, K]`` bf16 CUDA tensors whose rows are 16-byte
    aligned (row strides may exceed K).  N and K are arbitrary; when N is not a multiple of 8 the
    result is a ``[M, N]`` view of a row-padded buffer (see ``padded_empty``)."""
    assert _rows_ok(a) and _rows_ok(b) and a.dtype == b.dtype, "gemm_tn operands need 16-byte aligned rows"
    M, K = a.shape
    N, K2 = b.shape
    assert K == K2
    tf = a.dtype == torch.float32
    if out is None:
        out = padded_empty(M, N, torch.float32 if tf else out_dtype, a.device)
    assert out.stride(1) == 1
    mode = 2 if accumulate else _OUT[out.dtype]
    native.call("ceb_gemm_tf32_tn_ld" if tf else "ceb_gemm_bf16_tn_ld", native.ptr(a), a.stride(0), native.ptr(b),
                b.stride(0), native.ptr(out), out.stride(0), M, N, K, native.ptr(bias), 1 if relu else 0, mode,
                native.stream_ptr(a.device))
    return out


def gemm_tn_gelu(a: torch.Tensor, b: torch.Tensor, bias=None):
    """``pre = a @ b.T + bias`` and ``act = gelu_tanh(pre)`` from one GEMM (bf16, N % 32 == 0): returns
    ``(pre, act)``; the pre-activation is what the backward needs."""
    assert _rows_ok(a) and _rows_ok(b) and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    M, K = a.shape
    N = b.shape[0]
    assert b.shape[1] == K and N % 32 == 0
    pre = torch.empty((M, N), dtype=torch.bfloat16, device=a.device)
    act = torch.empty_like(pre)
    native.call("ceb_gemm_bf16_tn_gelu", native.ptr(a), a.stride(0), native.ptr(b), b.stride(0), native.ptr(pre),
                native.ptr(act), None, N, M, N, K, native.ptr(bias), 2, native.stream_ptr(a.device))
    return pre, act


def gemm_tn_gelu_grad(a: torch.Tensor, b: torch.Tensor, pre: torch.Tensor) -> torch.Tensor:
    """``(a @ b.T) * gelu_tanh'(pre)``: the GELU backward applied in the epilogue of the data-gradient GEMM."""
    assert _rows_ok(a) and _rows_ok(b) and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    M, K = a.shape
    N = b.shape[0]
    assert b.shape[1] == K and N % 32 == 0 and pre.shape == (M, N) and pre.is_contiguous() and pre.dtype == torch.bfloat16
    out = torch.empty((M, N), dtype=torch.bfloat16, device=a.device)
    native.call("ceb_gemm_bf16_tn_gelu", native.ptr(a), a.stride(0), native.ptr(b), b.stride(0), native.ptr(out), None,
                native.ptr(pre), N, M, N, K, None, 3, native.stream_ptr(a.device))
    return out


def gemm_nn(a: torch.Tensor, b: torch.Tensor, bias=None, relu=False, out_dtype=torch.bfloat16) -> torch.Tensor:
    """``act(a @ b + bias)`` with a ``[M, K]``, b ``[K, N]`` contiguous bf16 CUDA tensors."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.is_contiguous() and b.is_contiguous()
    M, K = a.shape
    K2, N = b.shape
    assert K == K2
    out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    native.call("ceb_gemm_bf16_nn", native.ptr(a), native.ptr(b), native.ptr(out), M, N, K, native.ptr(bias),
                1 if relu else 0, _OUT[out_dtype], native.stream_ptr(a.device))
    return out


def conv3x3_nhwc(x: torch.Tensor, w: torch.Tensor, relu=False, out_dtype=torch.bfloat16, relu_mask=None) -> torch.Tensor:
    """3x3 / stride 1 / pad 1 convolution.  ``x``: ``[N, H, W, Cin]`` bf16 contiguous, ``w``:
    ``[Cout, 3, 3, Cin]`` bf16 contiguous; returns ``[N, H, W, Cout]``.  ``relu_mask`` (same shape
    as the result): zero the result where the mask is <= 0 (ReLU backward fused into a dgrad)."""
    assert x.dtype == w.dtype and x.is_contiguous() and w.is_contiguous()
    n, h, wd, cin = x.shape
    cout = w.shape[0]
    if x.dtype == torch.float32:
        y = torch.empty((n, h, wd, cout), dtype=torch.float32, device=x.device)
        native.call("ceb_conv3x3_tf32_nhwc", native.ptr(x), native.ptr(w), native.ptr(y), n, h, wd, cin, cout, None,
                    1 if relu else 0, native.ptr(relu_mask), None, None, native.stream_ptr(x.device))
        return y
    assert x.dtype == torch.bfloat16
    y = torch.empty((n, h, wd, cout), dtype=out_dtype, device=x.device)
    native.call("ceb_conv3x3_bf16_nhwc", native.ptr(x), native.ptr(w), native.ptr(y), n, h, wd, cin, cout, None,
                1 if relu else 0, _OUT[out_dtype], native.ptr(relu_mask), native.stream_ptr(x.device))
    return y


def conv3x3_nhwc_res(x, w, resid, relu=False, relu_mask=None, keep_pre_add=False):
    """``act(conv3x3(x, w)) + resid`` in one kernel (bf16).  ``keep_pre_add``: returns
    ``(act(conv), act(conv) + resid)`` -- the first is what a later ReLU backward masks with."""
    assert x.dtype == w.dtype and x.is_contiguous() and w.is_contiguous()
    n, h, wd, cin = x.shape
    cout = w.shape[0]
    assert resid.shape == (n, h, wd, cout) and resid.dtype == x.dtype and resid.is_contiguous()
    y = torch.empty((n, h, wd, cout), dtype=x.dtype, device=x.device)
    y2 = torch.empty_like(y) if keep_pre_add else None
    if x.dtype == torch.float32:
        native.call("ceb_conv3x3_tf32_nhwc", native.ptr(x), native.ptr(w), native.ptr(y), n, h, wd, cin, cout, None,
                    1 if relu else 0, native.ptr(relu_mask), native.ptr(resid), native.ptr(y2),
                    native.stream_ptr(x.device))
        return (y, y2) if keep_pre_add else y
    native.call("ceb_conv3x3_bf16_nhwc_res", native.ptr(x), native.ptr(w), native.ptr(y), n, h, wd, cin, cout, None,
                1 if relu else 0, native.ptr(relu_mask), native.ptr(resid), native.ptr(y2), native.stream_ptr(x.device))
    return (y, y2) if keep_pre_add else y


def dgrad_weights(w_krsc_f32: torch.Tensor, dtype=torch.bfloat16) -> torch.Tensor:
    """``[Cin][3][3][Cout]`` weights of the data-gradient convolution (rotated 180 deg, in/out
    swapped) from fp32 ``[Cout][3][3][Cin]``."""
    if dtype == torch.float32:
        return prepare_conv_weights([w_krsc_f32], want_cast=False, want_rot=True, dtype=torch.float32)[1][0]
    cout, _, _, cin = w_krsc_f32.shape
    out = torch.empty((cin, 3, 3, cout), dtype=torch.bfloat16, device=w_krsc_f32.device)
    native.call("ceb_conv_w_dgrad_layout", native.ptr(w_krsc_f32), native.ptr(out), cout, cin,
                native.stream_ptr(out.device))
    return out


def prepare_conv_weights(weights_krsc_f32, want_cast=True, want_rot=True, dtype=torch.bfloat16):
    """Forward copies (bf16 casts; with ``dtype=float32`` the fp32 weights are used IN PLACE when they are
    16-byte aligned in the flat buffer, else copied) and data-gradient layouts (``dgrad_weights``) of a
    list of fp32 ``[Cout,3,3,Cin]`` conv weights in ONE launch.  Returns ``(casts, rots)`` (lists; None
    entries where not requested)."""
    import ctypes
    n = len(weights_krsc_f32)
    dev = weights_krsc_f32[0].device
    if dtype == torch.float32:
        inplace = [w.data_ptr() % 16 == 0 for w in weights_krsc_f32]
        casts = [None if (not want_cast or ok) else torch.empty(w.shape, dtype=dtype, device=dev)
                 for w, ok in zip(weights_krsc_f32, inplace)]
    else:
        casts = [torch.empty(w.shape, dtype=dtype, device=dev) if want_cast else None for w in weights_krsc_f32]
    rots = [torch.empty((w.shape[3], 3, 3, w.shape[0]), dtype=dtype, device=dev) if want_rot else None
            for w in weights_krsc_f32]
    PtrArr = ctypes.c_void_p * n
    IntArr = ctypes.c_int * n
    src = PtrArr(*[w.data_ptr() for w in weights_krsc_f32])
    cst = PtrArr(*[(c.data_ptr() if c is not None else None) for c in casts])
    rot = PtrArr(*[(r.data_ptr() if r is not None else None) for r in rots])
    cout = IntArr(*[int(w.shape[0]) for w in weights_krsc_f32])
    cin = IntArr(*[int(w.shape[3]) for w in weights_krsc_f32])
    if dtype == torch.float32 and not want_rot and all(c is None for c in casts):
        return list(weights_krsc_f32) if want_cast else casts, rots          # nothing to launch
    native.call("ceb_conv_weights_prepare_f32" if dtype == torch.float32 else "ceb_conv_weights_prepare", n,
                ctypes.cast(src, ctypes.c_void_p), ctypes.cast(cst, ctypes.c_void_p),
                ctypes.cast(rot, ctypes.c_void_p), ctypes.cast(cout, ctypes.c_void_p), ctypes.cast(cin, ctypes.c_void_p),
                native.stream_ptr(dev))
    if dtype == torch.float32 and want_cast:
        casts = [c if c is not None else w for c, w in zip(casts, weights_krsc_f32)]
    return casts, rots


def cast_bf16(src: torch.Tensor) -> torch.Tensor:
    out = torch.empty(src.shape, dtype=torch.bfloat16, device=src.device)
    native.call("ceb_cast_f32_bf16", native.ptr(src), native.ptr(out), src.numel(), native.stream_ptr(src.device))
    return out


def maxpool_nhwc(y: torch.Tensor, k: int) -> torch.Tensor:
    n, h, w, c = y.shape
    out = torch.empty((n, h // k, w // k, c), dtype=y.dtype, device=y.device)
    native.call("ceb_maxpool_nhwc_f32" if y.dtype == torch.float32 else "ceb_maxpool_nhwc_bf16", native.ptr(y), native.ptr(out), n, h, w, c, k, native.stream_ptr(y.device))
    return out


def maxpool_bwd_nhwc(y: torch.Tensor, dp: torch.Tensor, k: int, relu_mask: bool) -> torch.Tensor:
    n, h, w, c = y.shape
    dz = torch.empty_like(y)
    assert dp.dtype == y.dtype and dp.is_contiguous()
    native.call("ceb_maxpool_bwd_nhwc_f32" if y.dtype == torch.float32 else "ceb_maxpool_bwd_nhwc_bf16",
                native.ptr(y), native.ptr(dp), native.ptr(dz), n, h, w, c, k,
                1 if relu_mask else 0, native.stream_ptr(y.device))
    return dz


def gemm_mn_acc(a: torch.Tensor, b: torch.Tensor, c: torch.Tensor) -> torch.Tensor:
    """``c[M,N] (fp32) += a[K,M].T @ b[K,N]`` (bf16 operands with 16-byte aligned rows; the reduction
    runs over rows).  N % 64 == 0; M arbitrary."""
    assert _rows_ok(a) and _rows_ok(b) and a.dtype == b.dtype, "gemm_mn_acc operands need 16-byte aligned rows"
    assert c.dtype == torch.float32 and c.stride(1) == 1
    K, M = a.shape
    N = b.shape[1]
    native.call("ceb_gemm_tf32_mn_acc_ld" if a.dtype == torch.float32 else "ceb_gemm_bf16_mn_acc_ld", native.ptr(a), a.stride(0), native.ptr(b), b.stride(0), native.ptr(c),
                M, N, K, c.stride(0), native.stream_ptr(a.device))
    return c


def cast_transpose_bf16(src: torch.Tensor) -> torch.Tensor:
    """bf16 ``[C, R]`` transposed copy (rows padded to 16 bytes) of a contiguous fp32 ``[R, C]`` matrix."""
    assert src.dim() == 2 and src.dtype == torch.float32 and src.is_contiguous()
    R, C = src.shape
    out = padded_empty(C, R, torch.bfloat16, src.device)
    native.call("ceb_cast_transpose_f32_bf16", native.ptr(src), native.ptr(out), R, C, out.stride(0),
                native.stream_ptr(src.device))
    return out


def colsum_acc(x: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    """``out (fp32 [cols]) += x.sum(0)`` for bf16 ``x`` with 16-byte aligned rows and cols % 8 == 0."""
    assert _rows_ok(x) and x.shape[1] % 8 == 0 and out.dtype == torch.float32 and out.is_contiguous()
    native.call("ceb_colsum_bf16", native.ptr(x), x.stride(0), x.shape[0], x.shape[1], native.ptr(out),
                native.stream_ptr(x.device))
    return out


def im2col_c3(x_nhwc_f32: torch.Tensor, dtype=torch.bfloat16) -> torch.Tensor:
    """im2col matrix of the 3x3 / pad 1 windows of a 1- or 3-channel NHWC fp32 batch: ``[N*H*W, 64]`` bf16 or
    ``[N*H*W, 32]`` fp32 (9*Cin valid columns, zero padded to one 128-byte row)."""
    n, h, w, c = x_nhwc_f32.shape
    assert c in (1, 3) and x_nhwc_f32.is_contiguous() and x_nhwc_f32.dtype == torch.float32
    f32 = dtype == torch.float32
    out = torch.empty((n * h * w, 32 if f32 else 64), dtype=dtype, device=x_nhwc_f32.device)
    native.call("ceb_im2col3x3", native.ptr(x_nhwc_f32), native.ptr(out), n, h, w, c, 1 if f32 else 0,
                native.stream_ptr(out.device))
    return out


def _stem_weight_pad(w_krsc_f32: torch.Tensor, dtype=torch.bfloat16) -> torch.Tensor:
    """``[Cout][3][3][Cin]`` fp32 -> ``[Cout][64]`` bf16 / ``[Cout][32]`` fp32 (K-major B operand, zero padded)."""
    cout = w_krsc_f32.shape[0]
    k = w_krsc_f32[0].numel()
    pad = torch.zeros((cout, 32 if dtype == torch.float32 else 64), dtype=dtype, device=w_krsc_f32.device)
    pad[:, :k] = w_krsc_f32.reshape(cout, k)
    return pad


def relu_bwd(y: torch.Tensor, dy: torch.Tensor) -> torch.Tensor:
    dz = torch.empty_like(y)
    assert dy.dtype == y.dtype
    native.call("ceb_relu_bwd_f32" if y.dtype == torch.float32 else "ceb_relu_bwd_bf16", native.ptr(y),
                native.ptr(dy), native.ptr(dz), y.numel(),
                native.stream_ptr(y.device))
    return dz


def conv3x3_wgrad_nhwc(dz: torch.Tensor, x: torch.Tensor, dw: torch.Tensor) -> torch.Tensor:
    """``dw[Cout,3,3,Cin] (fp32, contiguous) += wgrad(dz[N,H,W,Cout], x[N,H,W,Cin])`` (bf16 inputs)."""
    assert dz.dtype == x.dtype and dz.is_contiguous() and x.is_contiguous()
    assert dw.dtype == torch.float32 and dw.is_contiguous()
    n, h, wd, cin = x.shape
    cout = dz.shape[3]
    native.call("ceb_conv3x3_wgrad_tf32_nhwc" if x.dtype == torch.float32 else "ceb_conv3x3_wgrad_bf16_nhwc", native.ptr(dz), native.ptr(x), native.ptr(dw), n, h, wd, cin, cout,
                native.stream_ptr(x.device))
    return dw


def conv2d_nhwc(x, w, ks=3, stride=1, pad=1, relu=False, relu_mask=None, resid=None, keep_pre_add=False):
    """General NHWC convolution on the tile kernel: ``x [N,H,W,Cin]``, ``w [Cout,ks,ks,Cin]`` (both bf16, or both
    fp32 for the TF32 path), ks in {1, 3}, stride in {1, 2}, pad in {0, 1}; any map whose OUTPUT width is <= 128.
    Optional fused ReLU, ReLU-backward mask and residual add as in ``conv3x3_nhwc(_res)``."""
    assert x.dtype == w.dtype and x.is_contiguous() and w.is_contiguous()
    n, h, wd, cin = x.shape
    cout = w.shape[0]
    ho, wo = (h + 2 * pad - ks) // stride + 1, (wd + 2 * pad - ks) // stride + 1
    y = torch.empty((n, ho, wo, cout), dtype=x.dtype, device=x.device)
    y2 = torch.empty_like(y) if (resid is not None and keep_pre_add) else None
    native.call("ceb_conv2d_nhwc", native.ptr(x), native.ptr(w), native.ptr(y), n, h, wd, cin, cout, ks, stride, pad,
                1 if relu else 0, 1 if x.dtype == torch.float32 else 0, native.ptr(relu_mask), native.ptr(resid),
                native.ptr(y2), native.stream_ptr(x.device))
    return (y, y2) if y2 is not None else y


def conv3x3_wgrad_strided(dz, x, dw, stride):
    """``dw[Cout,3,3,Cin] (fp32) +=`` weight gradient of a 3x3 / pad 1 / stride 1|2 convolution: ``x`` is the INPUT
    ``[N,H,W,Cin]``, ``dz`` the output gradient ``[N,Ho,Wo,Cout]`` (any map size with Wo <= 64 bf16 / 32 fp32)."""
    assert dz.dtype == x.dtype and dz.is_contiguous() and x.is_contiguous() and dw.dtype == torch.float32 and dw.is_contiguous()
    n, h, wd, cin = x.shape
    native.call("ceb_conv3x3_wgrad_nhwc", native.ptr(dz), native.ptr(x), native.ptr(dw), n, h, wd, cin, dz.shape[3],
                stride, 1 if x.dtype == torch.float32 else 0, native.stream_ptr(x.device))
    return dw


def subsample2(x):
    """``x[:, ::2, ::2, :]`` of an NHWC tensor as a dense tensor."""
    n, h, w, c = x.shape
    out = torch.empty((n, (h + 1) // 2, (w + 1) // 2, c), dtype=x.dtype, device=x.device)
    native.call("ceb_subsample2_nhwc", native.ptr(x), native.ptr(out), n, h, w, c * x.element_size(),
                native.stream_ptr(x.device))
    return out


def upsample2_zero(g, h, w):
    """``up[:, ::2, ::2, :] = g``, zero elsewhere, on an ``h x w`` grid."""
    n, ho, wo, c = g.shape
    up = torch.empty((n, h, w, c), dtype=g.dtype, device=g.device)
    native.call("ceb_upsample2_zero_nhwc", native.ptr(g), native.ptr(up), n, h, w, ho, wo, c * g.element_size(),
                native.stream_ptr(g.device))
    return up


def conv_supported(x, weight, stride, padding) -> bool:
    """3x3 / pad 1 and 1x1 / pad 0 convolutions with stride 1 or 2 and 64-multiple channels on maps up to 64
    pixels wide (every layer of ResNet-9 / ResNet-18 / FixupResNet18 except the 3- or 1-channel stems, which
    are lowered to im2col + GEMM by the whole-network paths), when the weight lives NHWC in the flat buffer."""
    if not (x.is_cuda and x.dim() == 4 and weight.dim() == 4):
        return False
    co, ci, kh, kw = weight.shape
    if (kh, kw, padding) not in ((3, 3, 1), (1, 1, 0)) or stride not in (1, 2) or ci % 64 or co % 64:
        return False
    h, w = x.shape[2], x.shape[3]
    wo = (w + 2 * padding - kw) // stride + 1
    kmax = 32 if act_dtype() == torch.float32 else 64
    if wo < 1 or wo > kmax or w > 128:
        return False
    return weight.permute(0, 2, 3, 1).is_contiguous()


class _ConvFn(torch.autograd.Function):
    """One convolution (3x3 / pad 1 or 1x1 / pad 0, stride 1 or 2, + optional ReLU) on the tcgen05 kernels.
    Activations cross the boundary as channels_last NCHW tensors (= NHWC buffers).  Backward:
      * data gradient: stride 1 -> the transposed-weight convolution; stride 2 -> the output gradient is spread
        onto the input grid with zeros (``upsample2_zero``) and then convolved at stride 1 (3x3), or is a GEMM
        followed by that spreading (1x1);
      * weight gradient: the MN-major wgrad kernel, whose TMA box steps through the input with the stride
        (3x3), or the plain MN-major GEMM over the sub-sampled input (1x1)."""

    @staticmethod
    def forward(ctx, x, weight, relu, stride):
        dt = act_dtype()
        ks = weight.shape[2]
        pad = 1 if ks == 3 else 0
        xn = x.permute(0, 2, 3, 1)
        xb = (xn if xn.dtype == dt else xn.to(dt)).contiguous()
        wk = weight.permute(0, 2, 3, 1)
        if dt == torch.float32:
            wb = wk.float().contiguous()
            if wb.data_ptr() % 16:
                wb = wb.clone()
        else:
            wb = cast_bf16(wk) if wk.dtype == torch.float32 else wk.to(torch.bfloat16).contiguous()
        y = conv2d_nhwc(xb, wb, ks, stride, pad, relu=relu)
        ctx.save_for_backward(xb, y if relu else xb)
        ctx.weight, ctx.relu, ctx.x_dtype, ctx.geom = weight, relu, x.dtype, (ks, stride, pad)
        return y.permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        xb, y = ctx.saved_tensors
        weight = ctx.weight
        ks, stride, pad = ctx.geom
        n, h, w, cin = xb.shape
        dz = dy.permute(0, 2, 3, 1).to(xb.dtype).contiguous()
        if ctx.relu:
            dz = relu_bwd(y, dz)
        cout = dz.shape[3]
        dx = dw = None
        if ctx.needs_input_grad[1]:
            target = weight.grad if (isinstance(weight, torch.nn.Parameter) and weight.grad is not None
                                     and weight.grad.permute(0, 2, 3, 1).is_contiguous()) else None
            acc = target.permute(0, 2, 3, 1) if target is not None else \
                torch.zeros(weight.permute(0, 2, 3, 1).shape, dtype=torch.float32, device=dz.device)
            if ks == 3:
                conv3x3_wgrad_strided(dz, xb, acc, stride)
            else:
                xs = subsample2(xb) if stride == 2 else xb
                gemm_mn_acc(dz.view(-1, cout), xs.view(-1, cin), acc.view(cout, cin))
            if target is None:
                dw = acc.permute(0, 3, 1, 2)
        if ctx.needs_input_grad[0]:
            wk = weight.data.permute(0, 2, 3, 1)
            wk = wk if wk.dtype == torch.float32 else wk.float().contiguous()
            if ks == 3:
                wd = dgrad_weights(wk, dtype=xb.dtype)
                g = upsample2_zero(dz, h, w) if stride == 2 else dz
                dxn = conv2d_nhwc(g, wd, 3, 1, 1)
            else:
                wt = _w_rows(wk.reshape(cout, cin), True, xb.dtype)                 # [Cin, Cout], K-major
                g = gemm_tn(dz.view(-1, cout), wt).view(n, dz.shape[1], dz.shape[2], cin)
                dxn = upsample2_zero(g.contiguous(), h, w) if stride == 2 else g
            dx = dxn.permute(0, 3, 1, 2).to(ctx.x_dtype)
        return dx, dw, None, None


def stem_supported(x, weight, stride, padding) -> bool:
    """The 1- or 3-channel first layer (3x3 / s1 / p1): lowered to im2col + the tensor-core GEMMs."""
    if not (x.is_cuda and x.dim() == 4 and weight.dim() == 4) or x.requires_grad:
        return False
    co, ci, kh, kw = weight.shape
    return ((kh, kw, stride, padding) == (3, 3, 1, 1) and ci in (1, 3) and co % 64 == 0 and x.shape[1] == ci
            and weight.permute(0, 2, 3, 1).is_contiguous())


class _StemFn(torch.autograd.Function):
    """First convolution of an image model (K = 9 or 27 is too thin for an implicit-GEMM tile): one pass writes
    the im2col matrix (zero padded to one 128-byte row), the layer then IS a K = 64 / 32 GEMM; the weight
    gradient is the MN-major GEMM over the same matrix.  The input needs no gradient."""

    @staticmethod
    def forward(ctx, x, weight, relu):
        dt = act_dtype()
        xb = x.float().permute(0, 2, 3, 1).contiguous()
        col = im2col_c3(xb, dt)
        wk = weight.permute(0, 2, 3, 1)
        y = gemm_tn(col, _stem_weight_pad(wk.float(), dt), relu=relu)
        n, _, h, w = x.shape
        ctx.save_for_backward(col, y if relu else col)
        ctx.weight, ctx.relu = weight, relu
        return y.view(n, h, w, -1).permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dy):
        col, y = ctx.saved_tensors
        weight = ctx.weight
        co = weight.shape[0]
        dz = dy.permute(0, 2, 3, 1).to(col.dtype).contiguous().view(-1, co)
        if ctx.relu:
            dz = relu_bwd(y, dz)
        k = weight[0].numel()
        g0 = torch.zeros((co, 64), dtype=torch.float32, device=dz.device)
        gemm_mn_acc(dz, col, g0[:, :col.shape[1]])
        target = weight.grad if (isinstance(weight, torch.nn.Parameter) and weight.grad is not None
                                 and weight.grad.permute(0, 2, 3, 1).is_contiguous()) else None
        if target is not None:
            target.permute(0, 2, 3, 1).reshape(co, k).add_(g0[:, :k])
            return None, None, None
        return None, g0[:, :k].reshape(co, weight.shape[2], weight.shape[3], weight.shape[1]).permute(0, 3, 1, 2), None


def conv2d(x, weight, bias, stride, padding, relu=False):
    if weight.shape[1] in (1, 3):
        y = _StemFn.apply(x, weight, relu)
        return y if bias is None else y + bias.view(1, -1, 1, 1)
    y = _ConvFn.apply(x, weight, relu, stride)
    if bias is not None:
        y = y + bias.view(1, -1, 1, 1)
    return y


def _weight_storage(weight: torch.Tensor):
    """``(storage [rows, cols] contiguous, transposed?)``: an ``nn.Linear`` weight ``[N, K]`` is its own
    storage; a GPT-2 ``Conv1D`` passes ``weight.t()`` of an ``[K, N]`` parameter."""
    if weight.is_contiguous():
        return weight, False
    if weight.t().is_contiguous():
        return weight.t(), True
    return None, False


def linear_supported(x, weight) -> bool:
    if not (x.is_cuda and weight.dim() == 2 and x.shape[-1] == weight.shape[1]):
        return False
    n, k = weight.shape
    store, transposed = _weight_storage(weight)
    if store is None or store.dtype != torch.float32 or k % 64 != 0 or n < 64:
        return False
    return n % 64 == 0 or not transposed        # Conv1D storage: N is the wgrad GEMM's 64-multiple side


def _as_rows(t: torch.Tensor, dtype=None) -> torch.Tensor:
    """2-D operand (``dtype``, default: the active precision's) with 16-byte aligned rows, without a copy
    when it already is one."""
    dtype = dtype or act_dtype()
    t2 = t.reshape(-1, t.shape[-1])
    if t2.dtype == dtype and _rows_ok(t2):
        return t2
    out = padded_empty(t2.shape[0], t2.shape[1], dtype, t2.device)
    out.copy_(t2)
    return out


def _w_rows(store: torch.Tensor, transpose: bool, dtype) -> torch.Tensor:
    """K-major GEMM operand made from an fp32 weight ``store`` (optionally transposed): a bf16 cast, or for
    the TF32 path the fp32 weight itself (no copy unless it must be transposed or re-aligned)."""
    if dtype == torch.bfloat16:
        return cast_transpose_bf16(store) if transpose else cast_bf16(store)
    if transpose:
        R, C = store.shape
        out = padded_empty(C, R, torch.float32, store.device)
        out.copy_(store.t())
        return out
    return store if _rows_ok(store) else _as_rows(store, torch.float32)


def _weight_grad(xb, dyb, weight, transposed):
    """``dW += dY^T X`` (split-K MN-major GEMM, fp32 accumulate) -- straight into ``weight.grad`` when that exists
    (the flat gradient buffer), in which case ``None`` is returned."""
    store, _ = _weight_storage(weight)
    param = weight._base if (transposed and weight._base is not None) else weight
    target = param.grad if (isinstance(param, torch.nn.Parameter) and param.grad is not None
                            and param.grad.is_contiguous()) else None
    acc = target if target is not None else torch.zeros_like(store, dtype=torch.float32)
    if transposed:
        gemm_mn_acc(xb, dyb, acc)          # dW_storage[K, N] += X^T dY
    else:
        gemm_mn_acc(dyb, xb, acc)          # dW[N, K] += dY^T X
    if target is not None:
        return None
    return acc.t() if transposed else acc


def _bias_grad(dyb, bias):
    """Column sums of the output gradient, accumulated straight into ``bias.grad`` when that exists."""
    if dyb.shape[1] % 8 == 0 and dyb.dtype == torch.bfloat16:
        direct = (isinstance(bias, torch.nn.Parameter) and bias.grad is not None
                  and bias.grad.is_contiguous() and bias.grad.dtype == torch.float32)
        acc_b = bias.grad if direct else torch.zeros(dyb.shape[1], dtype=torch.float32, device=dyb.device)
        colsum_acc(dyb, acc_b)
        return None if direct else acc_b
    return dyb.float().sum(0)


class _LinearFn(torch.autograd.Function):
    """``y = x W^T + b`` on the tcgen05 GEMMs.  Both GEMMs that read the weight want it K-major, so the
    fp32 master weight is cast to bf16 once per direction: as stored for one of them, through the
    transposing cast for the other (``nn.Linear`` ``[N, K]``: plain forward / transposed dgrad; GPT-2
    ``Conv1D`` ``[K, N]``: the other way round).  The weight gradient is the MN-major split-K GEMM
    accumulating in fp32 -- directly into ``weight.grad`` when that already exists (the flat gradient
    buffer) -- and the bias gradient a column-sum kernel over the bf16 output gradient.  N need not be
    a multiple of 8: outputs / gradients then live in row-padded buffers (LM head, V = 50 262)."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        store, transposed = _weight_storage(weight)
        xb = _as_rows(x)
        wb = _w_rows(store, transposed, xb.dtype)                                 # [N, K]
        bf = bias.float().contiguous() if bias is not None else None
        y = gemm_tn(xb, wb, bias=bf)
        ctx.save_for_backward(xb)
        ctx.transposed = transposed
        ctx.weight = weight
        ctx.bias = bias
        ctx.x_shape = x.shape
        ctx.x_dtype = x.dtype
        return y.view(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        (xb,) = ctx.saved_tensors
        weight, transposed, bias = ctx.weight, ctx.transposed, ctx.bias
        store, _ = _weight_storage(weight)
        dyb = _as_rows(dy, xb.dtype)
        dx = None
        if ctx.needs_input_grad[0]:
            wt = _w_rows(store, not transposed, xb.dtype)                        # [K, N]
            dx = gemm_tn(dyb, wt)
            dx = dx.view(ctx.x_shape).to(ctx.x_dtype)
        dw = _weight_grad(xb, dyb, weight, transposed) if ctx.needs_input_grad[1] else None
        db = _bias_grad(dyb, bias) if (bias is not None and ctx.needs_input_grad[2]) else None
        return dx, dw, db


def linear(x, weight, bias=None):
    return _LinearFn.apply(x, weight, bias)


class _MLPGeluFn(torch.autograd.Function):
    """The transformer MLP ``proj(gelu_tanh(fc(x)))`` with the activation inside the GEMM epilogues: the fc GEMM
    writes the pre-activation and its GELU in one pass, and the data-gradient GEMM of proj multiplies by
    ``GELU'(pre)`` while storing -- the ``[rows, 4E]`` tensors are touched by no element-wise kernel at all
    (reference: pytorch_transformers' MLP called from /root/reference/CommEfficient/gpt2_train.py:259-262)."""

    @staticmethod
    def forward(ctx, x, w_fc, b_fc, w_proj, b_proj):
        s_fc, t_fc = _weight_storage(w_fc)
        s_pr, t_pr = _weight_storage(w_proj)
        xb = _as_rows(x, torch.bfloat16)
        pre, act = gemm_tn_gelu(xb, _w_rows(s_fc, t_fc, torch.bfloat16),
                                bias=b_fc.float().contiguous() if b_fc is not None else None)
        y = gemm_tn(act, _w_rows(s_pr, t_pr, torch.bfloat16), bias=b_proj.float().contiguous() if b_proj is not None else None)
        ctx.save_for_backward(xb, pre, act)
        ctx.params = (w_fc, b_fc, w_proj, b_proj, t_fc, t_pr)
        ctx.x_shape, ctx.x_dtype = x.shape, x.dtype
        return y.view(*x.shape[:-1], w_proj.shape[0])

    @staticmethod
    def backward(ctx, dy):
        xb, pre, act = ctx.saved_tensors
        w_fc, b_fc, w_proj, b_proj, t_fc, t_pr = ctx.params
        s_fc, _ = _weight_storage(w_fc)
        s_pr, _ = _weight_storage(w_proj)
        dyb = _as_rows(dy, torch.bfloat16)
        need = ctx.needs_input_grad
        d_pre = gemm_tn_gelu_grad(dyb, _w_rows(s_pr, not t_pr, torch.bfloat16), pre)
        dw_pr = _weight_grad(act, dyb, w_proj, t_pr) if need[3] else None
        db_pr = _bias_grad(dyb, b_proj) if (b_proj is not None and need[4]) else None
        dx = None
        if need[0]:
            dx = gemm_tn(d_pre, _w_rows(s_fc, not t_fc, torch.bfloat16)).view(ctx.x_shape).to(ctx.x_dtype)
        dw_fc = _weight_grad(xb, d_pre, w_fc, t_fc) if need[1] else None
        db_fc = _bias_grad(d_pre, b_fc) if (b_fc is not None and need[2]) else None
        return dx, dw_fc, db_fc, dw_pr, db_pr


def mlp_gelu_supported(x, w_fc, w_proj) -> bool:
    return (linear_supported(x, w_fc) and w_fc.shape[0] % 64 == 0 and w_proj.dim() == 2
            and w_proj.shape[1] == w_fc.shape[0] and _weight_storage(w_proj)[0] is not None
            and w_proj.dtype == torch.float32 and w_proj.shape[0] % 64 == 0)


def mlp_gelu(x, w_fc, b_fc, w_proj, b_proj):
    return _MLPGeluFn.apply(x, w_fc, b_fc, w_proj, b_proj)


# ---------------------------------------------------------------------------- LM head + cross-entropy
def lm_head_supported(hidden2d, weight) -> bool:
    return (hidden2d.is_cuda and hidden2d.dim() == 2 and weight.dim() == 2 and weight.is_contiguous()
            and weight.dtype == torch.float32 and weight.shape[1] == hidden2d.shape[1]
            and weight.shape[1] % 64 == 0 and weight.shape[0] >= 64)


class _LMHeadCE(torch.autograd.Function):
    """Per-row cross-entropy of ``hidden @ W^T`` against ``labels`` (0 where ``labels == ignore``),
    computing logits ONLY for the rows whose label is not ignored.  In PersonaChat-style batches
    just the reply tokens of the gold candidate carry a label (reference:
    /root/reference/CommEfficient/gpt2_train.py:76-99 runs the full ``[N*C*T, 50 262]`` head and lets
    ``ignore_index=-1`` drop most of it); rows with an ignored label contribute neither loss nor
    gradient, so skipping them is exact.  Rows are compacted on the device (valid first, stable) and
    the three GEMMs plus the CE backward take the valid-row count from a DEVICE int, so nothing here
    synchronises with the host and the whole thing can sit inside a CUDA graph.  The logits buffer
    (row stride padded to 8) is overwritten in place by its own gradient."""

    @staticmethod
    def forward(ctx, hidden, weight, labels, ignore_index):
        from . import tfm  # noqa: F401  (registers the CE signatures)
        R, E = hidden.shape
        V = weight.shape[0]
        dev = hidden.device
        labels = labels.reshape(-1)
        valid = labels != ignore_index
        order = torch.argsort((~valid).to(torch.uint8), stable=True)
        nvalid = valid.sum().to(torch.int32).reshape(1)
        h_sel = hidden.detach().to(torch.bfloat16).index_select(0, order)
        lab_sel = labels.index_select(0, order).contiguous()
        wb = cast_bf16(weight.detach())
        logits = padded_empty(R, V, torch.bfloat16, dev)
        st = native.stream_ptr(dev)
        native.call("ceb_gemm_bf16_tn_rows", native.ptr(h_sel), h_sel.stride(0), native.ptr(wb), wb.stride(0),
                    native.ptr(logits), logits.stride(0), R, V, E, None, 0, 0, native.ptr(nvalid), st)
        loss_sel = torch.empty(R, dtype=torch.float32, device=dev)
        lse = torch.empty(R, dtype=torch.float32, device=dev)
        native.call("ceb_softmax_ce_fwd", native.ptr(logits), logits.stride(0), R, V, native.ptr(lab_sel),
                    int(ignore_index), native.ptr(loss_sel), native.ptr(lse), st)
        loss = torch.empty(R, dtype=torch.float32, device=dev).index_copy_(0, order, loss_sel)
        ctx.save_for_backward(h_sel, lab_sel, lse, logits, order, nvalid)
        ctx.weight = weight
        ctx.ignore_index = int(ignore_index)
        ctx.h_dtype = hidden.dtype
        return loss

    @staticmethod
    def backward(ctx, gloss):
        h_sel, lab_sel, lse, logits, order, nvalid = ctx.saved_tensors
        weight = ctx.weight
        R, E = h_sel.shape
        V = weight.shape[0]
        dev = h_sel.device
        st = native.stream_ptr(dev)
        g_sel = gloss.float().index_select(0, order).contiguous()
        native.call("ceb_softmax_ce_bwd_rows", native.ptr(logits), native.ptr(logits), logits.stride(0), R, V,
                    native.ptr(lab_sel), ctx.ignore_index, native.ptr(lse), native.ptr(g_sel), native.ptr(nvalid), st)
        dlogits = logits
        dh = None
        if ctx.needs_input_grad[0]:
            wt = cast_transpose_bf16(weight.detach())                       # [E, V], K-major for the dgrad
            # K = V = 50 262 with only (rows / 128) x (E / 256) output tiles: accumulate into fp32 so the kernel may
            # split the reduction over the idle SMs (out_mode 2), then round once
            dh32 = torch.zeros((R, E), dtype=torch.float32, device=dev)
            native.call("ceb_gemm_bf16_tn_rows", native.ptr(dlogits), dlogits.stride(0), native.ptr(wt), wt.stride(0),
                        native.ptr(dh32), dh32.stride(0), R, E, V, None, 0, 2, native.ptr(nvalid), st)
            dh_sel = cast_bf16(dh32)
            dh = torch.empty_like(dh_sel).index_copy_(0, order, dh_sel).to(ctx.h_dtype)
        dw = None
        if ctx.needs_input_grad[1]:
            target = weight.grad if (isinstance(weight, torch.nn.Parameter) and weight.grad is not None
                                     and weight.grad.is_contiguous()) else None
            acc = target if target is not None else torch.zeros_like(weight, dtype=torch.float32)
            native.call("ceb_gemm_bf16_mn_acc_rows", native.ptr(dlogits), dlogits.stride(0), native.ptr(h_sel),
                        h_sel.stride(0), native.ptr(acc), V, E, R, acc.stride(0), native.ptr(nvalid), st)
            dw = None if target is not None else acc
        return dh, dw, None, None


def lm_head_ce(hidden2d, weight, labels, ignore_index=-1):
    return _LMHeadCE.apply(hidden2d, weight, labels, ignore_index)


# ---------------------------------------------------------------------------- ResNet-9 on the native kernels
def _krsc(t: torch.Tensor) -> torch.Tensor:
    """fp32 ``[Cout,3,3,Cin]`` contiguous view of a conv weight / gradient stored NHWC in the flat buffer."""
    v = t.permute(0, 2, 3, 1)
    if not v.is_contiguous():
        raise RuntimeError("native conv path needs the flat buffer in 'nhwc' layout (--flat_layout nhwc)")
    return v


def resnet9_supported(model, x: torch.Tensor) -> bool:
    n = model.n
    # CIFAR (3 x 32 x 32) and FEMNIST (1 x 28 x 28: maps 28 -> 14 -> 7 -> 3, final 3 x 3 max-pool)
    if not x.is_cuda or x.dim() != 4 or x.shape[2] != x.shape[3] or x.shape[2] not in (32, 28):
        return False
    if n.prep.do_batchnorm:
        return False
    ch = n.channels
    if ch["prep"] != 64 or any(ch[k] % 64 for k in ("layer1", "layer2", "layer3")) or x.shape[1] not in (1, 3):
        return False
    fp = getattr(model, "_flat_params", None)
    return fp is not None and fp.layout == "nhwc"


def _r9_convs(net):
    return [net.prep.conv, net.layer1.conv, net.res1.res1.conv, net.res1.res2.conv, net.layer2.conv,
            net.layer3.conv, net.res3.res1.conv, net.res3.res2.conv]


def _r9_features(x, net):
    """Forward of the convolutional trunk: pooled bf16 features ``[N, C3]``, the activations the
    backward needs, and the data-gradient weight layouts (prepared in the same launch as the bf16
    forward copies)."""
    convs = _r9_convs(net)
    dt = act_dtype()
    # one launch: 7 casts + 7 dgrad layouts (TF32 path: the fp32 weights are read in place, 7 dgrad layouts)
    wb, wd = prepare_conv_weights([_krsc(c.weight.data) for c in convs[1:]], dtype=dt)
    xb = x.float().permute(0, 2, 3, 1).contiguous()           # NHWC fp32 (no copy for channels_last input)
    col = im2col_c3(xb, dt)                                   # [N*H*W, 64] bf16 / [N*H*W, 32] fp32
    a0 = gemm_tn(col, _stem_weight_pad(_krsc(convs[0].weight.data), dt), relu=True).view(x.shape[0], x.shape[2],
                                                                                         x.shape[3], -1)
    a1 = conv3x3_nhwc(a0, wb[0], relu=True)
    p1 = maxpool_nhwc(a1, 2)
    r1a = conv3x3_nhwc(p1, wb[1], relu=True)
    r1b, s1 = conv3x3_nhwc_res(r1a, wb[2], p1, relu=True, keep_pre_add=True)      # s1 = p1 + r1b in the epilogue
    a2 = conv3x3_nhwc(s1, wb[3], relu=True)
    p2 = maxpool_nhwc(a2, 2)
    a3 = conv3x3_nhwc(p2, wb[4], relu=True)
    p3 = maxpool_nhwc(a3, 2)
    r3a = conv3x3_nhwc(p3, wb[5], relu=True)
    r3b, s3 = conv3x3_nhwc_res(r3a, wb[6], p3, relu=True, keep_pre_add=True)
    f = maxpool_nhwc(s3, s3.shape[1]).view(x.shape[0], -1)     # 4 x 4 (CIFAR) or 3 x 3 (FEMNIST) -> 1 x 1
    return f, (col, a0, a1, p1, r1a, r1b, s1, a2, p2, a3, p3, r3a, r3b, s3), [None] + wd


def _r9_backward(net, saved, wd, df, fp=None):
    """Backward of the trunk from ``df`` (bf16 ``[N, C3]``): weight gradients are accumulated into
    the flat fp32 gradient buffer."""
    col, a0, a1, p1, r1a, r1b, s1, a2, p2, a3, p3, r3a, r3b, s3 = saved
    convs = _r9_convs(net)
    gw = [_krsc(c.weight.grad) for c in convs]                       # fp32 KRSC views of the flat gradient
    n = a0.shape[0]
    ds3 = maxpool_bwd_nhwc(s3, df.view(n, 1, 1, -1), s3.shape[1], relu_mask=False)

    dz = relu_bwd(r3b, ds3)
    conv3x3_wgrad_nhwc(dz, r3a, gw[7])
    dz = conv3x3_nhwc(dz, wd[7], relu_mask=r3a)
    conv3x3_wgrad_nhwc(dz, p3, gw[6])
    if fp is not None:
        # res3 (72 % of all coordinates) and the head are final: hand them to whoever consumes the
        # gradient (the worker sketches them on a second stream under the rest of this backward)
        lo, hi = fp.range_of([convs[6].weight, convs[7].weight, net.linear.weight])
        fp.range_final(lo, hi)
    dp3 = conv3x3_nhwc_res(dz, wd[6], ds3)                # ds3 + dgrad in the epilogue

    dz = maxpool_bwd_nhwc(a3, dp3, 2, relu_mask=True)
    conv3x3_wgrad_nhwc(dz, p2, gw[5])
    dp2 = conv3x3_nhwc(dz, wd[5])

    dz = maxpool_bwd_nhwc(a2, dp2, 2, relu_mask=True)
    conv3x3_wgrad_nhwc(dz, s1, gw[4])
    ds1 = conv3x3_nhwc(dz, wd[4])

    dz = relu_bwd(r1b, ds1)
    conv3x3_wgrad_nhwc(dz, r1a, gw[3])
    dz = conv3x3_nhwc(dz, wd[3], relu_mask=r1a)
    conv3x3_wgrad_nhwc(dz, p1, gw[2])
    dp1 = conv3x3_nhwc_res(dz, wd[2], ds1)

    dz = maxpool_bwd_nhwc(a1, dp1, 2, relu_mask=True)
    conv3x3_wgrad_nhwc(dz, a0, gw[1])
    da0 = conv3x3_nhwc(dz, wd[1], relu_mask=a0)

    g0 = torch.zeros((da0.shape[3], 64), dtype=torch.float32, device=da0.device)
    gemm_mn_acc(da0.view(-1, da0.shape[3]), col, g0[:, :col.shape[1]])
    k0 = gw[0][0].numel()                                          # 27 (RGB) or 9 (grey) stem taps
    gw[0].view(da0.shape[3], k0).add_(g0[:, :k0])


class _ResNet9Fn(torch.autograd.Function):
    """Whole-network forward/backward of ResNet-9 (no BatchNorm) on the hand-written kernels:
    implicit-GEMM conv + fused ReLU (tcgen05), fused ReLU/max-pool backward, MN-major wgrad
    accumulating straight into the flat fp32 gradient buffer, and the 3-channel stem lowered to
    im2col + the same tensor-core GEMMs.  Returns logits; the 512 -> classes head (5 MFLOP) runs on
    library ops here (``_ResNet9CEFn`` fuses it with the loss)."""

    @staticmethod
    def forward(ctx, x, net, *weights):
        f, saved, wd = _r9_features(x, net)
        logits = (f.float() @ net.linear.weight.data.t()) * net.out_scale
        ctx.net, ctx.wd, ctx.saved, ctx.f = net, wd, saved, f
        ctx.fp = _FP_OF.get(id(net))
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        net, f = ctx.net, ctx.f
        dl = dlogits.float() * net.out_scale
        net.linear.weight.grad.add_(dl.t() @ f.float())
        df = (dl @ net.linear.weight.data).to(f.dtype).contiguous()
        _r9_backward(net, ctx.saved, ctx.wd, df, ctx.fp)
        return (None, None) + (None,) * (len(ctx.needs_input_grad) - 2)


native.register_signatures({
    "ceb_linear_ce_head": ([native.c_ptr, native.c_ptr, native.c_f32, native.c_ptr, native.c_i32, native.c_i32,
                            native.c_i32, native.c_ptr, native.c_ptr, native.c_ptr, native.c_ptr, native.c_ptr,
                            native.c_ptr], native.c_i32),
})


def linear_ce_head(feat, weight, scale, labels, g=None, dW=None):
    """Fused classifier head + cross-entropy (``ceb_linear_ce_head``).  ``g is None``: returns per-sample
    ``(loss, correct)``; otherwise the backward for ``sum(g * loss)``: returns ``dfeat`` (bf16) and
    accumulates into ``dW``."""
    N, F_ = feat.shape
    C = weight.shape[0]
    dev = feat.device
    labels = labels.contiguous()
    fn = "ceb_linear_ce_head_f32" if feat.dtype == torch.float32 else "ceb_linear_ce_head"
    if g is None:
        loss = torch.empty(N, dtype=torch.float32, device=dev)
        correct = torch.empty(N, dtype=torch.float32, device=dev)
        native.call(fn, native.ptr(feat), native.ptr(weight), float(scale), native.ptr(labels), N, F_,
                    C, None, native.ptr(loss), native.ptr(correct), None, None, native.stream_ptr(dev))
        return loss, correct
    dfeat = torch.empty_like(feat)
    native.call(fn, native.ptr(feat), native.ptr(weight), float(scale), native.ptr(labels), N, F_, C,
                native.ptr(g), None, None, native.ptr(dfeat), native.ptr(dW), native.stream_ptr(dev))
    return dfeat


def head_supported(net, feat_dim: int) -> bool:
    w = net.linear.weight
    return (w.shape[0] <= 128 and w.shape[0] * feat_dim * 4 <= 224 * 1024 and feat_dim % 64 == 0 and feat_dim <= 2048
            and w.is_contiguous()
            and w.dtype == torch.float32 and w.grad is not None and w.grad.is_contiguous())


class _ResNet9CEFn(torch.autograd.Function):
    """ResNet-9 trunk + classifier head + per-sample cross-entropy / top-1 correctness: the head, the
    scale, softmax, NLL, arg-max and all of their backward are ONE kernel each way."""

    @staticmethod
    def forward(ctx, x, labels, net, *weights):
        f, saved, wd = _r9_features(x, net)
        f = f.contiguous()
        loss, correct = linear_ce_head(f, net.linear.weight.data, net.out_scale, labels)
        ctx.net, ctx.wd, ctx.saved, ctx.f, ctx.labels = net, wd, saved, f, labels
        ctx.fp = _FP_OF.get(id(net))
        ctx.mark_non_differentiable(correct)
        return loss, correct

    @staticmethod
    def backward(ctx, gloss, _gcorrect):
        net = ctx.net
        g = gloss.float().contiguous()
        df = linear_ce_head(ctx.f, net.linear.weight.data, net.out_scale, ctx.labels, g=g, dW=net.linear.weight.grad)
        _r9_backward(net, ctx.saved, ctx.wd, df, ctx.fp)
        return (None, None, None) + (None,) * (len(ctx.needs_input_grad) - 3)


def resnet9_ce_per_sample(model, x: torch.Tensor, labels: torch.Tensor):
    net = model.n
    _FP_OF[id(net)] = getattr(model, "_flat_params", None)
    weights = [p for p in net.parameters() if p.requires_grad]
    return _ResNet9CEFn.apply(x, labels, net, *weights)


_FP_OF = {}          # id(inner net) -> FlatParams of the model it belongs to (set on every call)


def resnet9_forward(model, x: torch.Tensor) -> torch.Tensor:
    net = model.n
    _FP_OF[id(net)] = getattr(model, "_flat_params", None)
    weights = [p for p in net.parameters() if p.requires_grad]
    return _ResNet9Fn.apply(x, net, *weights)
