"""Exact top-k by magnitude.

CUDA tensors go through the native radix select (``csrc/select.cu``); CPU tensors
through the torch oracle.  Reference: ``_topk`` in
/root/reference/CommEfficient/utils.py:232-252.
"""
from __future__ import annotations

from typing import Tuple

import torch

from . import native
from .ref import topk_dense_ref, topk_sparse_ref


def topk_sparse(vec: torch.Tensor, k: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """``(idx int32[k], val float32[k])`` of the k largest-|v| entries (unordered)."""
    if vec.dim() != 1:
        raise ValueError("topk_sparse expects a 1-D tensor")
    k = min(int(k), vec.numel())
    if not vec.is_cuda:
        idx, val = topk_sparse_ref(vec, k)
        return idx.to(torch.int32), val
    v = native.require_f32_contig(vec, "vec")
    out_idx = torch.empty(k, dtype=torch.int32, device=v.device)
    out_val = torch.zeros(k, dtype=torch.float32, device=v.device)
    scratch = native.select_scratch(v.device)
    native.call("ceb_topk_select", native.ptr(v), v.numel(), k, native.ptr(scratch), scratch.numel(),
                native.ptr(out_idx), native.ptr(out_val), native.stream_ptr(v.device))
    return out_idx, out_val


def topk_dense(vec: torch.Tensor, k: int) -> torch.Tensor:
    """Dense tensor keeping only the k largest-|v| entries (each row for 2-D input)."""
    if not vec.is_cuda:
        return topk_dense_ref(vec, k)
    if vec.dim() == 2:
        return torch.stack([topk_dense(row.contiguous(), k) for row in vec])
    idx, val = topk_sparse(vec.contiguous().float(), k)
    out = torch.zeros_like(vec, dtype=torch.float32)
    out[idx.long()] = val
    return out.to(vec.dtype)
