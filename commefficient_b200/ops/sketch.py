"""Count-sketch operations on a device (native kernels) or CPU (torch oracle).

``SketchPlan`` owns the hash material for one ``(d, c, r, num_blocks)`` geometry
on one device; the functional ops take a plan plus raw ``(r, c)`` tables.  This is
the layer ``csvec.CSVec`` (API-compatible object) and the federated server build
on.  Reference call sites: /root/reference/CommEfficient/fed_worker.py:312-320,
fed_aggregator.py:464-467,568-613.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from . import native, ref
from .hashing import SketchHashes, make_hashes


class SketchPlan:
    _cache: Dict[Tuple, "SketchPlan"] = {}

    def __init__(self, d: int, c: int, r: int, num_blocks: int, device):
        self.hashes: SketchHashes = make_hashes(d, c, r, num_blocks)
        self.d, self.c, self.r, self.num_blocks = int(d), int(c), int(r), int(num_blocks)
        self.n_tokens = self.hashes.n_tokens
        self.device = torch.device(device)
        self.is_cuda = self.device.type == "cuda"
        if self.is_cuda:
            h = self.hashes
            self.packed = h.packed.to(self.device)
            blk = h.block_offsets.to(torch.int32)
            neg = h.block_signs < 0
            blk = torch.where(neg, blk | torch.tensor(-2 ** 31, dtype=torch.int32), blk)
            self.blk = blk.contiguous().to(self.device)
            self._est = None
            self._l2_bufs = {}              # per stream: (row sums of squares, [scale, l2]) -- see l2_scale

    @classmethod
    def get(cls, d, c, r, num_blocks, device) -> "SketchPlan":
        device = torch.device(device)
        if device.type == "cuda" and device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        key = (int(d), int(c), int(r), int(num_blocks), str(device))
        plan = cls._cache.get(key)
        if plan is None:
            plan = cls(d, c, r, num_blocks, device)
            cls._cache[key] = plan
        return plan

    def new_table(self) -> torch.Tensor:
        return torch.zeros(self.r, self.c, dtype=torch.float32, device=self.device)

    def est_buffer(self) -> torch.Tensor:
        if self._est is None:
            self._est = torch.empty(self.d, dtype=torch.float32, device=self.device)
        return self._est


def accumulate(plan: SketchPlan, table: torch.Tensor, vec: torch.Tensor, scale: float = 1.0,
               weights: Optional[torch.Tensor] = None, wd: float = 0.0) -> None:
    """``table += S(scale * (vec + wd * weights))``."""
    if vec.numel() != plan.d:
        raise ValueError("vector has %d entries, sketch expects %d" % (vec.numel(), plan.d))
    if not plan.is_cuda:
        v = vec if weights is None or wd == 0.0 else vec + wd * weights
        ref.sketch_accum_ref(table, v, plan.hashes, scale)
        return
    native.require_f32_contig(vec, "vec")
    native.require_f32_contig(table, "table")
    w = None
    if weights is not None and wd != 0.0:
        w = native.require_f32_contig(weights, "weights")
    native.call("ceb_sketch_accumulate", native.ptr(vec), native.ptr(w), float(wd), float(scale), plan.d,
                native.ptr(plan.packed), plan.n_tokens, native.ptr(plan.blk), plan.num_blocks,
                native.ptr(table), plan.r, plan.c, native.stream_ptr(plan.device))


def accumulate_range(plan: SketchPlan, table: torch.Tensor, vec: torch.Tensor, lo: int, hi: int, scale: float = 1.0,
                     weights: Optional[torch.Tensor] = None, wd: float = 0.0) -> None:
    """``table += S(scale * (vec + wd * weights))`` for coordinates ``lo <= i < hi`` only (the sum over a
    partition of ``[0, d)`` equals ``accumulate``)."""
    lo, hi = int(lo), int(hi)
    if not (0 <= lo <= hi <= plan.d):
        raise ValueError("bad coordinate range [%d, %d) for d=%d" % (lo, hi, plan.d))
    if hi == lo:
        return
    if not plan.is_cuda:
        v = vec if weights is None or wd == 0.0 else vec + wd * weights
        part = torch.zeros_like(v)
        part[lo:hi] = v[lo:hi]
        ref.sketch_accum_ref(table, part, plan.hashes, scale)
        return
    native.require_f32_contig(vec, "vec")
    native.require_f32_contig(table, "table")
    w = None
    if weights is not None and wd != 0.0:
        w = native.require_f32_contig(weights, "weights")
    native.call("ceb_sketch_accumulate_range", native.ptr(vec), native.ptr(w), float(wd), float(scale), plan.d, lo, hi,
                native.ptr(plan.packed), plan.n_tokens, native.ptr(plan.blk), plan.num_blocks, native.ptr(table),
                plan.r, plan.c, native.stream_ptr(vec.device))


def l2_scale(plan: SketchPlan, table: torch.Tensor, clip: float = 0.0) -> torch.Tensor:
    """Device tensor ``[scale, l2]`` with ``l2 = sqrt(median_r sum_c T^2)`` and
    ``scale = clip / l2`` if ``l2 >= clip > 0`` else 1 -- no host sync."""
    if not plan.is_cuda:
        l2 = ref.l2estimate_ref(table)
        s = clip / l2 if (clip > 0 and l2 >= clip) else 1.0
        return torch.tensor([s, l2], dtype=torch.float32)
    # the result stays on the device and is consumed by a later kernel of the same stream: one buffer pair per
    # stream, so clients simulated concurrently (fed_worker.py: client lanes) cannot overwrite each other's scale
    key = torch.cuda.current_stream(plan.device).cuda_stream
    bufs = plan._l2_bufs.get(key)
    if bufs is None:
        bufs = (torch.zeros(plan.r, dtype=torch.float32, device=plan.device),
                torch.zeros(2, dtype=torch.float32, device=plan.device))
        plan._l2_bufs[key] = bufs
    native.call("ceb_sketch_l2_scale", native.ptr(table), plan.r, plan.c, float(clip),
                native.ptr(bufs[0]), native.ptr(bufs[1]), native.stream_ptr(plan.device))
    return bufs[1]


def axpy(y: torch.Tensor, x: torch.Tensor, alpha: float = 1.0, dev_scale: Optional[torch.Tensor] = None,
         zero_x: bool = False) -> None:
    """``y += alpha * dev_scale[0] * x`` (optionally clearing ``x``)."""
    if not y.is_cuda:
        s = alpha * (float(dev_scale[0]) if dev_scale is not None else 1.0)
        y.add_(x, alpha=s)
        if zero_x:
            x.zero_()
        return
    native.call("ceb_axpy_dev", native.ptr(y), native.ptr(x), y.numel(), float(alpha), native.ptr(dev_scale),
                1 if zero_x else 0, native.stream_ptr(y.device))


def server_update(Vv: torch.Tensor, Ve: Optional[torch.Tensor], g: torch.Tensor, inv_n: float, rho: float,
                  error_type: str) -> None:
    """``Vv = rho*Vv + g*inv_n``; virtual: ``Ve += Vv``; local: ``Ve = Vv``."""
    mode = {"none": 0, "local": 1, "virtual": 2}[error_type]
    if not Vv.is_cuda:
        Vv.mul_(rho).add_(g, alpha=inv_n)
        if mode == 2:
            Ve.add_(Vv)
        elif mode == 1:
            Ve.copy_(Vv)
        return
    native.call("ceb_sketch_server_update", native.ptr(Vv), native.ptr(Ve if mode else None), native.ptr(g),
                float(inv_n), float(rho), mode, Vv.numel(), native.stream_ptr(Vv.device))


def unsketch_topk(plan: SketchPlan, table: torch.Tensor, k: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """Heavy-hitter recovery: ``(idx int32[k], val float32[k])`` of the k largest
    median estimates."""
    k = min(int(k), plan.d)
    if not plan.is_cuda:
        idx, val = ref.unsketch_ref(table, plan.hashes, k)
        return idx.to(torch.int32), val
    native.require_f32_contig(table, "table")
    out_idx = torch.empty(k, dtype=torch.int32, device=plan.device)
    out_val = torch.zeros(k, dtype=torch.float32, device=plan.device)
    scratch = native.select_scratch(plan.device)
    native.call("ceb_sketch_unsketch_topk", native.ptr(table), native.ptr(plan.packed), plan.n_tokens,
                native.ptr(plan.blk), plan.num_blocks, plan.r, plan.c, plan.d, k,
                native.ptr(plan.est_buffer()), native.ptr(scratch), scratch.numel(),
                native.ptr(out_idx), native.ptr(out_val), native.stream_ptr(plan.device))
    return out_idx, out_val


def query(plan: SketchPlan, table: torch.Tensor, idx: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Median estimates of the given coordinates (all coordinates when ``idx`` is None)."""
    if not plan.is_cuda:
        est = ref.sketch_estimate_ref(table, plan.hashes)
        return est if idx is None else est[idx.long()]
    count = plan.d if idx is None else idx.numel()
    out = torch.empty(count, dtype=torch.float32, device=plan.device)
    if idx is not None:
        idx = idx.to(torch.int32).contiguous()
    native.call("ceb_sketch_query", native.ptr(table), native.ptr(plan.packed), plan.n_tokens,
                native.ptr(plan.blk), plan.num_blocks, plan.r, plan.c, plan.d, native.ptr(idx), count,
                native.ptr(out), native.stream_ptr(plan.device))
    return out


def apply_zero(plan: SketchPlan, idx: torch.Tensor, val: torch.Tensor, Vv: Optional[torch.Tensor],
               Ve: Optional[torch.Tensor], weights: Optional[torch.Tensor], lr=0.0,
               changed: Optional[torch.Tensor] = None) -> None:
    """Zero the buckets of the selected coordinates in ``Vv``/``Ve`` and apply
    ``weights[idx] -= lr * val`` (``lr`` float or length-d tensor)."""
    lr_dev = lr if (torch.is_tensor(lr) and lr.numel() == 1 and plan.is_cuda) else None   # device scalar
    if lr_dev is not None:
        native.call("ceb_sketch_apply_zero_dev", native.ptr(idx), native.ptr(val), idx.numel(),
                    native.ptr(plan.packed), plan.n_tokens, native.ptr(plan.blk), plan.r, plan.c,
                    native.ptr(Vv), native.ptr(Ve), native.ptr(weights), native.ptr(lr_dev), native.ptr(changed),
                    native.stream_ptr(plan.device))
        return
    lr_vec = lr if torch.is_tensor(lr) else None
    if not plan.is_cuda:
        ref.sketch_zero_buckets_ref([Vv, Ve], idx, val, plan.hashes)
        if weights is not None:
            li = idx.long()
            step = (lr_vec[li] if lr_vec is not None else lr) * val
            weights[li] -= step
            if changed is not None:
                changed[li[step != 0]] = 1
        return
    native.call("ceb_sketch_apply_zero", native.ptr(idx), native.ptr(val), idx.numel(),
                native.ptr(plan.packed), plan.n_tokens, native.ptr(plan.blk), plan.r, plan.c,
                native.ptr(Vv), native.ptr(Ve), native.ptr(weights),
                0.0 if lr_vec is not None else float(lr), native.ptr(lr_vec), native.ptr(changed),
                native.stream_ptr(plan.device))
