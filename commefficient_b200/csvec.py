"""``CSVec`` -- count sketch of a length-d vector, API-compatible with the external
``csvec`` package the reference imports (it is not part of the reference tree;
call sites: /root/reference/CommEfficient/fed_aggregator.py:464-467,585-595,
fed_worker.py:313-320, utils.py:305-313; semantics restated in SURVEY.md 2.2).

The object is a thin wrapper: hash material lives in a cached ``SketchPlan`` and
the ``(r, c)`` fp32 ``table`` is a plain tensor, so server code can hand tables to
the fused native kernels without going through this class.
"""
from __future__ import annotations

import copy
from typing import Optional

import numpy as np
import torch

from .ops import sketch as sk
from .ops.topk import topk_sparse


class CSVec:
    def __init__(self, d, c, r, doInitialize=True, device=None, numBlocks=1):
        self.d, self.c, self.r, self.numBlocks = int(d), int(c), int(r), int(numBlocks)
        if device is None:
            device = "cuda" if torch.cuda.is_available() else "cpu"
        self.device = torch.device(device)
        self.plan = sk.SketchPlan.get(self.d, self.c, self.r, self.numBlocks, self.device)
        self.table = self.plan.new_table() if doInitialize else None

    # -- structure ---------------------------------------------------------------
    def _same_geometry(self, other: "CSVec") -> bool:
        return (self.d, self.c, self.r, self.numBlocks) == (other.d, other.c, other.r, other.numBlocks) \
            and self.plan.device == other.plan.device

    def zero(self):
        self.table.zero_()

    def cpu_(self):
        self.device = torch.device("cpu")
        self.plan = sk.SketchPlan.get(self.d, self.c, self.r, self.numBlocks, self.device)
        self.table = self.table.cpu()

    def cuda_(self, device="cuda"):
        self.device = torch.device(device)
        self.plan = sk.SketchPlan.get(self.d, self.c, self.r, self.numBlocks, self.device)
        self.table = self.table.to(self.plan.device)

    def half_(self):
        raise NotImplementedError("sketch tables stay fp32 (atomics and medians are fp32)")

    def __deepcopy__(self, memo):
        new = CSVec(self.d, self.c, self.r, doInitialize=False, device=self.device, numBlocks=self.numBlocks)
        new.table = copy.deepcopy(self.table, memo)
        return new

    # -- arithmetic (linear, so all of it is table arithmetic) --------------------
    def __imul__(self, other):
        if isinstance(other, (int, float)):
            self.table.mul_(other)
            return self
        raise ValueError("can only multiply a CSVec by a scalar")

    def __truediv__(self, other):
        # in place, returns self -- the reference relies on this (utils.py:313)
        if isinstance(other, (int, float)):
            self.table.div_(other)
            return self
        raise ValueError("can only divide a CSVec by a scalar")

    def __iadd__(self, other):
        if isinstance(other, CSVec):
            self.accumulateCSVec(other)
            return self
        raise ValueError("can only add a CSVec to a CSVec")

    def __add__(self, other):
        out = copy.deepcopy(self)
        out += other
        return out

    # -- accumulation ------------------------------------------------------------
    def accumulateVec(self, vec: torch.Tensor):
        assert vec.dim() == 1 and vec.numel() == self.d, "vector length must equal d"
        v = vec.detach()
        if v.device != self.plan.device or v.dtype != torch.float32 or not v.is_contiguous():
            v = v.to(device=self.plan.device, dtype=torch.float32).contiguous()
        sk.accumulate(self.plan, self.table, v)

    def accumulateTable(self, table: torch.Tensor):
        if tuple(table.shape) != (self.r, self.c):
            raise ValueError("table shape %s does not match sketch (%d, %d)" % (tuple(table.shape), self.r, self.c))
        self.table += table.to(self.table.device)

    def accumulateCSVec(self, other: "CSVec"):
        assert self._same_geometry(other), "sketch geometries differ"
        self.table += other.table

    # -- queries -----------------------------------------------------------------
    def l2estimate(self) -> float:
        return float(sk.l2_scale(self.plan, self.table, 0.0)[1])

    def _findAllValues(self) -> torch.Tensor:
        return sk.query(self.plan, self.table)

    def unSketch(self, k: Optional[int] = None, epsilon: Optional[float] = None) -> torch.Tensor:
        """Dense length-d vector holding the recovered heavy hitters: the k largest
        median estimates, or (``epsilon``) every estimate with magnitude at least
        ``epsilon * l2estimate()``."""
        out = torch.zeros(self.d, dtype=torch.float32, device=self.plan.device)
        if k is not None:
            idx, val = sk.unsketch_topk(self.plan, self.table, k)
            out[idx.long()] = val
            return out
        if epsilon is None:
            raise ValueError("unSketch needs k or epsilon")
        est = self._findAllValues()
        if epsilon == 1:
            return est
        thr = epsilon * self.l2estimate()
        keep = est.abs() >= thr
        out[keep] = est[keep]
        return out

    def unSketchSparse(self, k: int):
        """``(idx int32[k], val float32[k])`` without materialising the dense vector."""
        return sk.unsketch_topk(self.plan, self.table, k)

    @classmethod
    def median(cls, csvecs):
        first = csvecs[0]
        for other in csvecs:
            assert first._same_geometry(other)
        out = copy.deepcopy(first)
        out.table = torch.median(torch.stack([c.table for c in csvecs]), dim=0)[0]
        return out


def topk(vec: torch.Tensor, k: int) -> torch.Tensor:
    """Dense top-k helper kept for symmetry with the upstream package."""
    idx, val = topk_sparse(vec, k)
    out = torch.zeros_like(vec)
    out[idx.long()] = val
    return out
