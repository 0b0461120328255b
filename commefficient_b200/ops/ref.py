"""Plain-PyTorch oracles for every native op.

These are (a) the numerics reference the CUDA kernels are tested against and
(b) the implementation used on CPU tensors (the CPU/gloo plumbing backend,
BASELINE.json config 1).  They restate the maths of the reference call sites:
``_topk`` (/root/reference/CommEfficient/utils.py:232-252), the CSVec methods the
reference uses and the per-mode server rules
(/root/reference/CommEfficient/fed_aggregator.py:483-613).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from .hashing import SketchHashes


# --------------------------------------------------------------------------- top-k

def topk_sparse_ref(vec: torch.Tensor, k: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """Indices and values of the k largest-magnitude entries of a 1-D tensor.
    NaNs sort as +inf magnitude in torch.topk; we map them to magnitude 0 so a
    NaN can never be selected (the reference works around NaN garbage by
    zero-initialising its outputs, utils.py:238-243)."""
    k = min(int(k), vec.numel())
    mag = vec.detach().float().abs()
    mag = torch.where(torch.isnan(mag), torch.zeros_like(mag), mag)
    if k == 0:
        return torch.zeros(0, dtype=torch.long, device=vec.device), vec[:0]
    # Deterministic rule shared with csrc/select.cu: everything above the k-th largest magnitude, then
    # the LOWEST-INDEX entries equal to it; output in (above: ascending index, ties: ascending index).
    thresh = torch.topk(mag, k, sorted=True).values[-1]
    above = torch.nonzero(mag > thresh).flatten()
    ties = torch.nonzero(mag == thresh).flatten()[: k - above.numel()]
    idx = torch.cat([above, ties])
    val = vec[idx]
    if float(thresh) == 0.0:
        val = torch.where(mag[idx] == 0, torch.zeros_like(val), val)     # NaN at magnitude 0 reads as 0
    return idx, val


def topk_dense_ref(vec: torch.Tensor, k: int) -> torch.Tensor:
    """Dense vector that keeps the k largest-magnitude entries (per row for 2-D)."""
    if vec.dim() == 1:
        out = torch.zeros_like(vec)
        idx, val = topk_sparse_ref(vec, k)
        out[idx] = val
        return out
    if vec.dim() == 2:
        return torch.stack([topk_dense_ref(row, k) for row in vec])
    raise ValueError("topk expects a 1-D or 2-D tensor")


# --------------------------------------------------------------------------- sketch

def sketch_accum_ref(table: torch.Tensor, vec: torch.Tensor, h: SketchHashes,
                     scale: float = 1.0) -> None:
    """table[r] += bincount(bucket_r, sign_r * vec * scale)."""
    v = vec.detach().to(torch.float32).cpu() * scale
    for row in range(h.r):
        b = h.coord_buckets(row)
        s = h.coord_signs(row)
        add = torch.bincount(b, weights=(s * v).double(), minlength=h.c).float()
        table[row] += add.to(table.device)


def sketch_estimate_ref(table: torch.Tensor, h: SketchHashes) -> torch.Tensor:
    """est[i] = median_r table[r, bucket_r(i)] * sign_r(i)  (lower median)."""
    t = table.detach().cpu()
    vals = torch.empty(h.r, h.d, dtype=torch.float32)
    for row in range(h.r):
        vals[row] = t[row, h.coord_buckets(row)] * h.coord_signs(row)
    return vals.median(dim=0)[0]


def unsketch_ref(table: torch.Tensor, h: SketchHashes, k: int) -> Tuple[torch.Tensor, torch.Tensor]:
    est = sketch_estimate_ref(table, h)
    idx, val = topk_sparse_ref(est, k)
    return idx.to(table.device), val.to(table.device)


def l2estimate_ref(table: torch.Tensor) -> float:
    return float(torch.sqrt(torch.median((table.double() ** 2).sum(dim=1))))


def sketch_zero_buckets_ref(tables, idx: torch.Tensor, val: torch.Tensor, h: SketchHashes) -> None:
    """Zero, in each given (r, c) table, every bucket touched by a selected
    coordinate whose value is non-zero (fed_aggregator.py:594-611; the reference
    re-sketches the update and zeroes ``nonzero()`` of that sketch, which differs
    only when two selected values cancel exactly inside one bucket)."""
    idx = idx.detach().cpu().long()
    keep = val.detach().cpu() != 0
    idx = idx[keep]
    if idx.numel() == 0:
        return
    n = h.n_tokens
    t, b = idx % n, idx // n
    for row in range(h.r):
        cols = (h.buckets[row, t] + h.block_offsets[b]) % h.c
        for tab in tables:
            if tab is not None:
                tab[row, cols.to(tab.device)] = 0


# --------------------------------------------------------------------------- server rules

def server_update_ref(mode: str, gradient: torch.Tensor, Vv: torch.Tensor, Ve: torch.Tensor,
                      *, rho: float, lr, k: int = 0, error_type: str = "none",
                      hashes: Optional[SketchHashes] = None, noise_std: float = 0.0,
                      generator=None):
    """Pure-torch server step.  Mutates ``Vv``/``Ve`` in place and returns the
    weight update either densely (``dense`` tensor) or sparsely
    (``(idx, val)``): ``(dense_or_None, sparse_or_None)`` with ``lr`` already
    applied.  ``lr`` is a float or a length-d tensor."""
    Vv.mul_(rho).add_(gradient)
    if mode == "fedavg":
        return Vv.clone(), None
    if mode == "uncompressed":
        upd = Vv
        if noise_std > 0:
            upd = Vv + torch.normal(0.0, noise_std, Vv.shape, generator=generator).to(Vv.device)
        return upd * lr, None
    if mode == "local_topk":
        return Vv * lr, None
    if mode == "true_topk":
        Ve.add_(Vv)
        idx, val = topk_sparse_ref(Ve, k)
        val = val.clone()
        Ve[idx] = 0
        Vv[idx] = 0
        lr_sel = lr[idx] if torch.is_tensor(lr) else lr
        return None, (idx, val * lr_sel)
    if mode == "sketch":
        if error_type == "local":
            Ve.copy_(Vv)
        elif error_type == "virtual":
            Ve.add_(Vv)
        idx, val = unsketch_ref(Ve, hashes, k)
        if error_type == "virtual":
            sketch_zero_buckets_ref([Ve, Vv], idx, val, hashes)
        else:
            sketch_zero_buckets_ref([Vv], idx, val, hashes)
            if error_type == "local":
                # Ve aliases Vv in the reference (fed_aggregator.py:580)
                Ve.copy_(Vv)
        lr_sel = lr[idx] if torch.is_tensor(lr) else lr
        return None, (idx, val * lr_sel)
    raise ValueError("unknown mode %r" % mode)
