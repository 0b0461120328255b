"""Dense-vector server rules, per-client state updates, clipping, DP noise and
byte-accounting helpers: native kernels on CUDA tensors (``csrc/dense.cu``), torch
on CPU tensors.  Reference: /root/reference/CommEfficient/fed_aggregator.py:483-566,
fed_worker.py:82-108,190-216,290-309.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import native
from .ref import topk_sparse_ref


def dense_server_update(Vv: torch.Tensor, g: torch.Tensor, inv_n: float, rho: float, weights: torch.Tensor,
                        lr, noise_std: float = 0.0, seed: int = 0, noise_offset: int = 0,
                        changed: Optional[torch.Tensor] = None) -> None:
    """``Vv = rho*Vv + g*inv_n``; ``weights -= lr * (Vv [+ N(0, noise_std^2)])``."""
    lr_vec = lr if torch.is_tensor(lr) else None
    if not Vv.is_cuda:
        Vv.mul_(rho).add_(g, alpha=inv_n)
        upd = Vv
        if noise_std > 0:
            gen = torch.Generator().manual_seed(seed + noise_offset)
            upd = Vv + torch.normal(0.0, noise_std, Vv.shape, generator=gen)
        step = upd * (lr_vec if lr_vec is not None else lr)
        weights.sub_(step)
        if changed is not None:
            changed[step != 0] = 1
        return
    native.call("ceb_dense_server_update", native.ptr(Vv), native.ptr(g), float(inv_n), float(rho),
                Vv.numel(), native.ptr(weights), 0.0 if lr_vec is not None else float(lr),
                native.ptr(lr_vec), float(noise_std), int(seed), int(noise_offset), native.ptr(changed),
                native.stream_ptr(Vv.device))


def true_topk_step(Vv: torch.Tensor, Ve: torch.Tensor, g: torch.Tensor, inv_n: float, rho: float, k: int,
                   weights: torch.Tensor, lr, changed: Optional[torch.Tensor] = None,
                   client_vel: Optional[torch.Tensor] = None,
                   vel_rows: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Momentum + virtual error + exact top-k + error/momentum zeroing + apply.
    Returns the selected ``(idx, val)`` (values before the learning rate)."""
    k = min(int(k), Vv.numel())
    lr_vec = lr if torch.is_tensor(lr) else None
    if not Vv.is_cuda:
        Vv.mul_(rho).add_(g, alpha=inv_n)
        Ve.add_(Vv)
        idx, val = topk_sparse_ref(Ve, k)
        val = val.clone()
        keep = val != 0
        li = idx[keep]
        Ve[li] = 0
        Vv[li] = 0
        step = val[keep] * (lr_vec[li] if lr_vec is not None else lr)
        weights[li] -= step
        if changed is not None:
            changed[li[step != 0]] = 1
        if client_vel is not None and vel_rows is not None and vel_rows.numel():
            client_vel[vel_rows.long().view(-1, 1), li.view(1, -1)] = 0
        return idx.to(torch.int32), val
    out_idx = torch.empty(k, dtype=torch.int32, device=Vv.device)
    out_val = torch.zeros(k, dtype=torch.float32, device=Vv.device)
    scratch = native.select_scratch(Vv.device)
    n_rows = 0 if vel_rows is None else int(vel_rows.numel())
    if vel_rows is not None:
        vel_rows = vel_rows.to(device=Vv.device, dtype=torch.int64).contiguous()
    native.call("ceb_true_topk_step", native.ptr(Vv), native.ptr(Ve), native.ptr(g), float(inv_n),
                float(rho), Vv.numel(), k, native.ptr(weights), 0.0 if lr_vec is not None else float(lr),
                native.ptr(lr_vec), native.ptr(changed), native.ptr(scratch), scratch.numel(),
                native.ptr(out_idx), native.ptr(out_val),
                native.ptr(client_vel if n_rows else None), native.ptr(vel_rows if n_rows else None),
                n_rows, native.stream_ptr(Vv.device))
    return out_idx, out_val


def client_step(grad: torch.Tensor, scale: float, vel: Optional[torch.Tensor], rho: float,
                err: Optional[torch.Tensor], acc: torch.Tensor, k: int = 0) -> None:
    """One client's local momentum / error feedback / (optional) local top-k, adding
    what the client transmits into ``acc``.  ``grad`` is scaled in place when there
    is neither velocity nor error.  ``vel``/``err`` are updated in place (rows of the
    device-resident client-state store)."""
    n = grad.numel()
    if not grad.is_cuda:
        x = grad.mul_(scale)
        if vel is not None:
            vel.mul_(rho).add_(x)
            x = vel
        if err is not None:
            err.add_(x)
            x = err
        if k and k > 0:
            idx, val = topk_sparse_ref(x, min(k, n))
            val = val.clone()
            acc[idx] += val
            nz = idx[val != 0]
            if err is not None:
                err[nz] = 0
            if vel is not None:
                vel[nz] = 0
        else:
            acc.add_(x)
        return
    k = min(int(k), n) if k and k > 0 else 0
    scratch = out_idx = out_val = None
    if k:
        scratch = native.select_scratch(grad.device)
        out_idx = torch.empty(k, dtype=torch.int32, device=grad.device)
        out_val = torch.zeros(k, dtype=torch.float32, device=grad.device)
    native.call("ceb_client_step", native.ptr(grad), float(scale), native.ptr(vel), float(rho),
                native.ptr(err), n, k, native.ptr(acc), native.ptr(scratch),
                0 if scratch is None else scratch.numel(), native.ptr(out_idx), native.ptr(out_val),
                native.stream_ptr(grad.device))


def scatter_add_zero(idx, val, acc=None, err=None, vel=None) -> None:
    if not idx.is_cuda:
        li = idx.long()
        nz = li[val != 0]
        if acc is not None:
            acc[li] += val
        if err is not None:
            err[nz] = 0
        if vel is not None:
            vel[nz] = 0
        return
    native.call("ceb_scatter_add_zero", native.ptr(idx), native.ptr(val), idx.numel(), native.ptr(acc),
                native.ptr(err), native.ptr(vel), native.stream_ptr(idx.device))


_sumsq_cache = {}


def clip_by_norm(x: torch.Tensor, clip: float, eps: float = 0.0) -> None:
    """In-place clip of ``x`` to L2 norm ``clip`` with no host sync.  ``eps > 0``
    follows ``torch.nn.utils.clip_grad_norm_`` (coef = min(1, clip/(norm+eps)));
    ``eps == 0`` follows the reference ``clip_grad`` (scale only when norm >= clip)."""
    if not x.is_cuda:
        norm = float(torch.linalg.vector_norm(x))
        if eps > 0:
            coef = min(1.0, clip / (norm + eps))
        else:
            coef = clip / norm if norm >= clip else 1.0
        if coef != 1.0:
            x.mul_(coef)
        return
    key = (x.device.index, torch.cuda.current_stream(x.device).cuda_stream)    # (concurrent client lanes)
    buf = _sumsq_cache.get(key)
    if buf is None:
        buf = torch.zeros(1, dtype=torch.float32, device=x.device)
        _sumsq_cache[key] = buf
    native.call("ceb_clip_by_norm", native.ptr(x), x.numel(), float(clip), float(eps), native.ptr(buf),
                native.stream_ptr(x.device))


def add_gaussian_noise(x: torch.Tensor, std: float, seed: int, offset: int = 0) -> None:
    if std == 0:
        return
    if not x.is_cuda:
        gen = torch.Generator().manual_seed((int(seed) * 1000003 + int(offset)) & 0x7FFFFFFFFFFFFFFF)
        x.add_(torch.normal(0.0, std, x.shape, generator=gen))
        return
    native.call("ceb_add_gaussian_noise", native.ptr(x), x.numel(), float(std), int(seed), int(offset),
                native.stream_ptr(x.device))


def fedavg_delta(acc: torch.Tensor, w0: torch.Tensor, wloc: torch.Tensor, weight: float) -> None:
    """``acc += weight * (w0 - wloc)``."""
    if not acc.is_cuda:
        acc.add_(w0 - wloc, alpha=weight)
        return
    native.call("ceb_fedavg_delta", native.ptr(acc), native.ptr(w0), native.ptr(wloc), float(weight),
                acc.numel(), native.stream_ptr(acc.device))


def count_changed(weights: torch.Tensor, prev: Optional[torch.Tensor], mask: Optional[torch.Tensor],
                  count_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``mask |= (weights != prev)``; ``prev = weights``; returns a 1-element int64
    tensor with ``mask.sum()`` (stays on the device)."""
    if not weights.is_cuda:
        if prev is not None:
            diff = weights != prev
            if mask is not None:
                mask |= diff.to(mask.dtype)
            prev.copy_(weights)
            total = (mask if mask is not None else diff).sum()
        else:
            total = mask.sum()
        return total.reshape(1).to(torch.int64)
    if count_out is None:
        count_out = torch.zeros(1, dtype=torch.int64, device=weights.device)
    native.call("ceb_count_changed", native.ptr(weights), native.ptr(prev), native.ptr(mask),
                weights.numel(), native.ptr(count_out), native.stream_ptr(weights.device))
    return count_out
