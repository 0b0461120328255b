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

import os

import torch
import torch.nn.functional as F

_native = False
_precision = "bf16"


def set_native(flag: bool) -> None:
    global _native
    _native = bool(flag)


def set_precision(precision: str) -> None:
    """``--precision``, honestly:
      * ``bf16``: bf16 activations, tcgen05 ``kind::f16`` MMAs, fp32 accumulation (the fast path);
      * ``tf32``: fp32 activations whose GEMM / convolution inputs the tensor cores read as TF32
        (``kind::tf32``) -- the arithmetic the reference's fp32 convolutions get from cuDNN, whose
        ``allow_tf32`` defaults to True; library fallbacks run with TF32 enabled as well;
      * ``fp32``: IEEE fp32 everywhere: no tensor-core layer runs (there is no fp32 MMA), the library path
        is used with TF32 switched OFF.
    The bf16-I/O transformer kernels (LayerNorm, GELU, cross-entropy, segment BatchNorm) only run under
    ``bf16``."""
    global _precision
    if precision not in ("bf16", "tf32", "fp32"):
        raise ValueError("unknown precision %r" % (precision,))
    _precision = precision
    if precision != "fp32":
        from . import tc
        tc.set_precision(precision)
    if torch.cuda.is_available():
        tf = precision != "fp32"
        torch.backends.cudnn.allow_tf32 = tf
        torch.backends.cuda.matmul.allow_tf32 = tf


def precision() -> str:
    return _precision


def native_enabled() -> bool:
    return _native and _precision != "fp32"


def _use_native(x: torch.Tensor) -> bool:
    return _native and x.is_cuda and _precision != "fp32"


def _use_native_bf16(x: torch.Tensor) -> bool:
    return _use_native(x) and _precision == "bf16"


def conv2d(x, weight, bias=None, stride=1, padding=0):
    if _use_native(x):
        from . import tc
        if tc.conv_supported(x, weight, stride, padding) or tc.stem_supported(x, weight, stride, padding):
            return tc.conv2d(x, weight, bias, stride, padding, relu=False)
    return F.conv2d(x, weight, bias, stride=stride, padding=padding)


def conv2d_relu(x, weight, bias=None, stride=1, padding=0):
    if _use_native(x):
        from . import tc
        if tc.conv_supported(x, weight, stride, padding) or tc.stem_supported(x, weight, stride, padding):
            return tc.conv2d(x, weight, bias, stride, padding, relu=True)
    return F.relu(F.conv2d(x, weight, bias, stride=stride, padding=padding))


def linear(x, weight, bias=None):
    if _use_native(x):
        from . import tc
        if tc.linear_supported(x, weight):
            return tc.linear(x, weight, bias)
    return F.linear(x, weight, bias)


def layer_norm(x, weight, bias, eps=1e-5):
    if _use_native_bf16(x) and weight.is_cuda:
        from . import tfm
        return tfm.layer_norm(x, weight, bias, eps)
    return F.layer_norm(x, (x.shape[-1],), weight, bias, eps)


def add_layer_norm(resid, branch, weight, bias, eps=1e-5, p_drop=0.0, salt=0):
    """``(resid + dropout(branch), LayerNorm(of that))`` -- the residual add (and the dropout in front of it)
    fused into the LayerNorm that follows (native path: one kernel each way, ``ops/tfm.py:_AddLayerNorm``;
    the mask is a counter hash of the position, regenerated in the backward)."""
    if (_use_native_bf16(resid) and weight.is_cuda and resid.dtype == torch.float32
            and resid.shape[-1] in (128, 256, 512, 768, 1024, 1280)):
        from . import tfm
        return tfm.add_layer_norm(resid, branch, weight, bias, eps, p_drop, salt,
                                  tfm.dropout_state(resid.device) if p_drop > 0 else None)
    if p_drop > 0:
        branch = F.dropout(branch, p_drop, training=True)
    new_resid = resid + branch
    return new_resid, layer_norm(new_resid, weight, bias, eps)


def fused_dropout_available(x: torch.Tensor) -> bool:
    """Whether the hooks below apply dropout inside their kernels (so the model must not apply it again)."""
    return _use_native_bf16(x)


def attention(qkv, B, T, n_head, p_drop=0.0, salt=0):
    """Causal multi-head self-attention over the fused projection ``qkv [B, T, 3E]`` -> ``[B, T, E]``
    (head dim 64).  Native path: tcgen05 flash-style kernels (``csrc/attn.cu``); library path: SDPA."""
    E = qkv.shape[-1] // 3
    if _use_native_bf16(qkv):
        from . import tfm
        if tfm.attention_supported(qkv, T, n_head):
            y = tfm.attention(qkv, B, T, n_head, p_drop, salt, tfm.dropout_state(qkv.device) if p_drop > 0 else None)
            return y.view(B, T, E)
    hd = E // n_head
    q, k, v = qkv.view(B, T, 3 * E).split(E, dim=2)
    q = q.view(B, T, n_head, hd).transpose(1, 2)
    k = k.view(B, T, n_head, hd).transpose(1, 2)
    v = v.view(B, T, n_head, hd).transpose(1, 2)
    y = F.scaled_dot_product_attention(q, k, v, is_causal=True, dropout_p=p_drop)
    return y.transpose(1, 2).reshape(B, T, E)


def embed(wte, wpe, input_ids, token_type_ids, p_drop=0.0, salt=0):
    """``dropout(wte[ids] + wpe[pos] + wte[token_types])`` -> fp32 ``[..., T, E]``."""
    T = input_ids.shape[-1]
    if _use_native_bf16(wte):
        from . import tfm
        if tfm.embed_supported(wte, wpe):
            return tfm.embed(wte, wpe, input_ids, token_type_ids, T, p_drop, salt,
                             tfm.dropout_state(wte.device) if p_drop > 0 else None)
    pos = torch.arange(T, device=input_ids.device)
    x = F.embedding(input_ids, wte) + F.embedding(pos, wpe)
    if token_type_ids is not None:
        x = x + F.embedding(token_type_ids, wte)
    return F.dropout(x, p_drop, training=True) if p_drop > 0 else x


def advance_dropout(device) -> None:
    """Once per training forward pass: move every native dropout site to a fresh mask."""
    if _native and _precision == "bf16" and torch.device(device).type == "cuda":
        from . import tfm
        tfm.dropout_state(device).advance()


_FUSE_MLP = os.environ.get("CEB_FUSE_MLP", "1") != "0"      # A/B switch for bench/ (the unfused chain is the fallback)


def mlp_gelu(x, w_fc, b_fc, w_proj, b_proj):
    """``proj(gelu_tanh(fc(x)))`` (weights in ``F.linear`` orientation ``[out, in]``).  Native bf16 path: the GELU and
    its backward live in the GEMM epilogues (``ops/tc.py:_MLPGeluFn``)."""
    if _use_native_bf16(x) and _FUSE_MLP:
        from . import tc
        if tc.mlp_gelu_supported(x, w_fc, w_proj):
            return tc.mlp_gelu(x, w_fc, b_fc, w_proj, b_proj)
    return linear(gelu_tanh(linear(x, w_fc, b_fc)), w_proj, b_proj)


def gelu_tanh(x):
    if _use_native_bf16(x) and x.numel() % 2 == 0:
        from . import tfm
        return tfm.gelu_tanh(x)
    return F.gelu(x, approximate="tanh")


def cross_entropy_rows(logits, labels, ignore_index=-1):
    """Per-row cross-entropy (0 where ``labels == ignore_index``) of ``[rows, V]`` logits."""
    if _use_native_bf16(logits) and logits.dtype == torch.bfloat16 and logits.dim() == 2 and logits.stride(1) == 1 \
            and logits.stride(0) % 2 == 0:
        from . import tfm
        return tfm.cross_entropy_rows(logits, labels, ignore_index)
    return F.cross_entropy(logits.float(), labels, ignore_index=ignore_index, reduction="none")


def lm_head_cross_entropy(hidden2d, weight, labels, ignore_index=-1):
    """Per-row cross-entropy of ``hidden2d @ weight.T`` (0 where ``labels == ignore_index``).  Native
    path: one fused function that only computes the rows with a label (``ops/tc.py:_LMHeadCE``)."""
    if _use_native_bf16(hidden2d):
        from . import tc
        if tc.lm_head_supported(hidden2d, weight):
            return tc.lm_head_ce(hidden2d, weight, labels, ignore_index)
    return cross_entropy_rows(linear(hidden2d, weight, None), labels, ignore_index)


# ---------------------------------------------------------------------------- segment-wise BatchNorm
# The reference runs every client's batch through the model separately, so BatchNorm statistics are
# per CLIENT batch (resnet9.py:17-29, fixup_resnet18.py:143-146).  To batch a rank's clients into one
# forward/backward (fed_worker._round_batched) BatchNorm must keep that semantics: statistics per
# (client, channel) over that client's samples only.  ``bn_segments(n)`` announces that the leading
# dimension of every activation is n equal-sized client segments; ``BatchNorm2d`` (a drop-in
# ``nn.BatchNorm2d``: same parameters, buffers and state_dict keys) then normalises per segment.
_bn_segments = 0            # number of client segments in the current batch (0: plain BatchNorm)
_bn_seg_ids = None          # [N] long: segment of every sample (None: equal-sized consecutive segments)
_bn_seg_counts = None       # [S] float: samples per segment


class bn_segments:
    """``with bn_segments(n):`` -- the batch is n equal-sized consecutive client segments;
    ``with bn_segments(n, seg=ids, counts=c):`` -- arbitrary sizes: ``ids[i]`` is the segment of sample i
    and ``c[s]`` the number of samples of segment s (device tensors, built once outside any graph capture)."""

    def __init__(self, n: int, seg=None, counts=None):
        self.n, self.seg = int(n), seg
        self.counts = counts.reshape(-1).float() if counts is not None else None

    def __enter__(self):
        global _bn_segments, _bn_seg_ids, _bn_seg_counts
        self.prev = (_bn_segments, _bn_seg_ids, _bn_seg_counts)
        _bn_segments, _bn_seg_ids, _bn_seg_counts = self.n, self.seg, self.counts
        return self

    def __exit__(self, *exc):
        global _bn_segments, _bn_seg_ids, _bn_seg_counts
        _bn_segments, _bn_seg_ids, _bn_seg_counts = self.prev
        return False


class _SegmentBNGeneral(torch.autograd.Function):
    """Segment-wise BatchNorm for segments of different sizes: per-sample partial sums are combined per
    segment with ``index_add_`` and the statistics gathered back per sample."""

    @staticmethod
    def forward(ctx, x, weight, bias, seg, counts, segments, eps):
        N, C = x.shape[0], x.shape[1]
        xf = x.reshape(N, C, -1).float()
        L = xf.shape[2]
        denom = (counts * L).clamp_min(1.0).view(segments, 1)
        mean = torch.zeros(segments, C, dtype=torch.float32, device=x.device).index_add_(0, seg, xf.sum(2)) / denom
        xc = xf - mean[seg].unsqueeze(2)
        var = torch.zeros(segments, C, dtype=torch.float32, device=x.device).index_add_(0, seg, (xc * xc).sum(2)) / denom
        rstd = torch.rsqrt(var + eps)
        xhat = xc * rstd[seg].unsqueeze(2)
        y = xhat * weight.float().view(1, C, 1) + bias.float().view(1, C, 1)
        ctx.save_for_backward(xhat.to(x.dtype), rstd, weight, seg, denom)
        ctx.shape, ctx.segments = x.shape, segments
        out = y.to(x.dtype).reshape(x.shape)
        if x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last):
            out = out.contiguous(memory_format=torch.channels_last)
        ctx.mark_non_differentiable(mean, var)
        return out, mean, var

    @staticmethod
    def backward(ctx, dy, _dmean, _dvar):
        xhat, rstd, weight, seg, denom = ctx.saved_tensors
        N, C, L = xhat.shape
        S = ctx.segments
        dyf = dy.reshape(N, C, L).float()
        xh = xhat.float()
        dbeta = dyf.sum(dim=(0, 2))
        dgamma = (dyf * xh).sum(dim=(0, 2))
        g = dyf * weight.float().view(1, C, 1)
        m1 = torch.zeros(S, C, dtype=torch.float32, device=dy.device).index_add_(0, seg, g.sum(2)) / denom
        m2 = torch.zeros(S, C, dtype=torch.float32, device=dy.device).index_add_(0, seg, (g * xh).sum(2)) / denom
        dx = (g - m1[seg].unsqueeze(2) - xh * m2[seg].unsqueeze(2)) * rstd[seg].unsqueeze(2)
        dx = dx.to(dy.dtype).reshape(ctx.shape)
        if dy.dim() == 4 and dy.is_contiguous(memory_format=torch.channels_last):
            dx = dx.contiguous(memory_format=torch.channels_last)
        return dx, dgamma.to(weight.dtype), dbeta.to(weight.dtype), None, None, None, None


class _SegmentBN(torch.autograd.Function):
    """y = (x - mean_seg) * rstd_seg * gamma + beta with statistics per (segment, channel).
    x: ``[S*n, C, H, W]`` (any memory format); fp32 statistics; two passes forward, two backward."""

    @staticmethod
    def forward(ctx, x, weight, bias, segments, eps):
        N, C = x.shape[0], x.shape[1]
        n = N // segments
        x5 = x.reshape(segments, n, C, -1)                               # [S, n, C, HW] (view when possible)
        xf = x5.float()
        var, mean = torch.var_mean(xf, dim=(1, 3), unbiased=False, keepdim=True)     # [S, 1, C, 1]
        rstd = torch.rsqrt(var + eps)
        xhat = (xf - mean) * rstd
        y = xhat * weight.float().view(1, 1, C, 1) + bias.float().view(1, 1, C, 1)
        ctx.save_for_backward(xhat.to(x.dtype), rstd, weight)
        ctx.shape = x.shape
        out = y.to(x.dtype).reshape(x.shape)
        if x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last):
            out = out.contiguous(memory_format=torch.channels_last)
        mean, var = mean.reshape(segments, C), var.reshape(segments, C)
        ctx.mark_non_differentiable(mean, var)
        return out, mean, var

    @staticmethod
    def backward(ctx, dy, _dmean, _dvar):
        xhat, rstd, weight = ctx.saved_tensors
        S, n, C, L = xhat.shape
        dyf = dy.reshape(S, n, C, L).float()
        xh = xhat.float()
        dbeta = dyf.sum(dim=(0, 1, 3))
        dgamma = (dyf * xh).sum(dim=(0, 1, 3))
        g = dyf * weight.float().view(1, 1, C, 1)
        m1 = g.mean(dim=(1, 3), keepdim=True)
        m2 = (g * xh).mean(dim=(1, 3), keepdim=True)
        dx = (g - m1 - xh * m2) * rstd
        dx = dx.to(dy.dtype).reshape(ctx.shape)
        if dy.dim() == 4 and dy.is_contiguous(memory_format=torch.channels_last):
            dx = dx.contiguous(memory_format=torch.channels_last)
        return dx, dgamma.to(weight.dtype), dbeta.to(weight.dtype), None, None


class BatchNorm2d(torch.nn.BatchNorm2d):
    """``nn.BatchNorm2d`` that, inside ``bn_segments(n)`` and in training mode, normalises each of the n
    equal client segments of the batch with its own statistics.  Running statistics are advanced once
    per call with the mean over the segments (the reference never federates them: SURVEY Q12)."""

    def forward(self, x):
        S = _bn_segments
        general = _bn_seg_ids is not None and _bn_seg_ids.shape[0] == x.shape[0]
        if not (self.training and S > 1 and self.affine and (general or x.shape[0] % S == 0)):
            return super().forward(x)
        native_ok = False
        if general and _use_native(x) and _precision in ("bf16", "tf32"):
            from . import tfm
            native_ok = tfm.segment_bn_supported(x, _bn_seg_ids, _bn_seg_counts)
        if native_ok:
            y, mean, var = tfm.segment_batch_norm(x, self.weight, self.bias, _bn_seg_ids, _bn_seg_counts, S, self.eps)
        elif general:
            y, mean, var = _SegmentBNGeneral.apply(x, self.weight, self.bias, _bn_seg_ids, _bn_seg_counts, S, self.eps)
        else:
            y, mean, var = _SegmentBN.apply(x, self.weight, self.bias, S, self.eps)
        if self.track_running_stats and self.running_mean is not None:
            with torch.no_grad():
                n = max(x.shape[0] // S, 1) * (x.numel() // (x.shape[0] * x.shape[1]))   # ~samples x pixels per segment
                m = self.momentum if self.momentum is not None else 0.1
                self.running_mean.mul_(1 - m).add_(mean.mean(0).to(self.running_mean.dtype), alpha=m)
                unbiased = var.mean(0) * (n / max(n - 1, 1))
                self.running_var.mul_(1 - m).add_(unbiased.to(self.running_var.dtype), alpha=m)
                self.num_batches_tracked += 1
        return y
