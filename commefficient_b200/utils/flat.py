"""Flat parameter / gradient buffers.

The reference flattens and un-flattens the model on every client
(``torch.cat`` of ~25-150 tensors and two launches per parameter,
/root/reference/CommEfficient/utils.py:254-297).  Here the trainable parameters
of a module *live* in one contiguous fp32 buffer and their ``.grad`` tensors in
a second one; the module's ``nn.Parameter`` objects are views.  Reading the
gradient vector, writing the weight vector, and zeroing gradients are then O(1)
launches, and the native kernels (sketch, top-k, apply) consume the buffers
directly.  The coordinate order is ``model.parameters()`` order restricted to
``requires_grad`` parameters -- identical to the reference, so checkpoints and
the sketch coordinate space agree.

The free functions at the bottom keep the reference's helper names for code
written against ``utils.get_param_vec`` & co.
"""
from __future__ import annotations

from typing import Tuple, List

import torch


class FlatParams:
    """Owns ``flat`` (weights) and ``grad`` (gradients) for one module replica."""

    def __init__(self, model: torch.nn.Module, device=None, align: int = 1, layout: str = "reference",
                 flat_alloc=None):
        """``layout="reference"``: every parameter occupies its slice in C-contiguous
        (out, in, kh, kw) order -- coordinate i of the flat vector is coordinate i of the
        reference's.  ``layout="nhwc"``: 4-D (convolution) weights are stored
        (out, kh, kw, in) -- the K-major "KRSC" layout tensor-core implicit GEMM wants -- and
        exposed to the module as channels_last views of the usual logical shape, so neither
        cuDNN nor the native kernels ever transpose weights or weight gradients.  Only the
        order of coordinates *inside* a conv weight changes (a relabelling the sketch and
        top-k are indifferent to); ``state_dict`` keys, shapes and values are unchanged."""
        if layout not in ("reference", "nhwc"):
            raise ValueError("layout must be 'reference' or 'nhwc'")
        self.layout = layout
        params = [p for p in model.parameters() if p.requires_grad]
        if not params:
            raise ValueError("model has no trainable parameters")
        self.model = model
        self.params: List[torch.nn.Parameter] = params
        self.shapes = [tuple(p.shape) for p in params]
        self.sizes = [p.numel() for p in params]
        self.offsets = []
        off = 0
        for n in self.sizes:
            self.offsets.append(off)
            off += n
        self.numel = off
        device = torch.device(device) if device is not None else params[0].device
        # Padded allocation so 16-byte vector accesses at the tail stay in bounds.
        pad = (-off) % 1024
        # ``flat_alloc(numel) -> tensor`` lets the weights live in peer-mapped (symmetric) memory
        self._flat_store = (flat_alloc(off + pad) if flat_alloc is not None
                            else torch.zeros(off + pad, dtype=torch.float32, device=device))
        self._grad_store = torch.zeros(off + pad, dtype=torch.float32, device=device)
        self.flat = self._flat_store[:off]
        self.grad = self._grad_store[:off]
        with torch.no_grad():
            for p, o, n, shp in zip(params, self.offsets, self.sizes, self.shapes):
                self._view(self.flat, o, n, shp).copy_(p.data.to(device=device, dtype=torch.float32))
        model.to(device)
        self._bind()

    def _view(self, store: torch.Tensor, o: int, n: int, shp):
        """Logical-shape view of one parameter's slice of ``store``."""
        flat = store[o:o + n]
        if self.layout == "nhwc" and len(shp) == 4:
            co, ci, kh, kw = shp
            return flat.view(co, kh, kw, ci).permute(0, 3, 1, 2)
        return flat.view(shp)

    def _bind(self):
        for p, o, n, shp in zip(self.params, self.offsets, self.sizes, self.shapes):
            p.data = self._view(self.flat, o, n, shp)
            p.grad = self._view(self.grad, o, n, shp)

    def rebind(self):
        """Re-establish the views (after ``model.to()`` / ``load_state_dict`` with
        ``assign=True`` or anything else that swapped ``p.data``)."""
        with torch.no_grad():
            for p, o, n, shp in zip(self.params, self.offsets, self.sizes, self.shapes):
                if p.data.data_ptr() != self.flat[o:o + n].data_ptr():
                    self._view(self.flat, o, n, shp).copy_(p.data)
        self._bind()

    # -- early hand-off of finished gradient ranges -------------------------------------
    # A consumer (the sketching worker) may set ``on_range_final`` to a callable ``(lo, hi)``; code that
    # knows a contiguous coordinate range of ``grad`` will not change any more during the current
    # backward pass calls ``range_final`` so the range can be consumed (sketched on a second stream)
    # while the rest of the backward is still running.
    on_range_final = None

    def range_of(self, params) -> Tuple[int, int]:
        """``(lo, hi)`` of the coordinates covered by ``params`` (which must be adjacent in the buffer)."""
        ids = {id(p) for p in params}
        idx = [i for i, p in enumerate(self.params) if id(p) in ids]
        if not idx:
            return 0, 0
        lo_i, hi_i = min(idx), max(idx)
        if hi_i - lo_i + 1 != len(idx):
            raise ValueError("parameters are not adjacent in the flat buffer")
        return self.offsets[lo_i], self.offsets[hi_i] + self.sizes[hi_i]

    def range_final(self, lo: int, hi: int) -> None:
        cb = self.on_range_final
        if cb is not None and hi > lo:
            cb(int(lo), int(hi))

    def zero_grad(self):
        self.grad.zero_()
        for p, o, n, shp in zip(self.params, self.offsets, self.sizes, self.shapes):
            if p.grad is None or p.grad.data_ptr() != self.grad[o:o + n].data_ptr():
                p.grad = self._view(self.grad, o, n, shp)

    def set_weights(self, vec: torch.Tensor):
        """``vec`` is in this buffer's own coordinate order."""
        self.flat.copy_(vec.reshape(-1))

    def to_reference_order(self, store: torch.Tensor) -> torch.Tensor:
        """Copy of ``store`` (weights- or gradient-shaped) in the reference's coordinate order."""
        if self.layout == "reference":
            return store.clone()
        return torch.cat([self._view(store, o, n, shp).reshape(-1)
                          for o, n, shp in zip(self.offsets, self.sizes, self.shapes)])

    def segment_vector(self, values) -> torch.Tensor:
        """Expand one scalar per parameter tensor into a length-d vector in flat
        order (used for per-group learning rates; fixes reference quirk Q9 where
        the LR vector was laid out in param-group order)."""
        out = torch.zeros(self._flat_store.numel(), dtype=torch.float32, device=self.flat.device)[: self.numel]
        for v, o, n in zip(values, self.offsets, self.sizes):
            out[o:o + n] = float(v)
        return out


# ---------------------------------------------------------------------------
# Reference-named helpers (work on any module, flat-backed or not)
# ---------------------------------------------------------------------------

def _trainable(model):
    return [p for p in model.parameters() if p.requires_grad]


def get_param_vec(model) -> torch.Tensor:
    """utils.py:281-286."""
    fp = getattr(model, "_flat_params", None)
    if fp is not None:
        return fp.flat.clone()
    return torch.cat([p.data.reshape(-1).float() for p in _trainable(model)])


def set_param_vec(model, vec: torch.Tensor) -> None:
    """utils.py:290-297."""
    fp = getattr(model, "_flat_params", None)
    if fp is not None:
        fp.set_weights(vec.to(fp.flat.device))
        return
    off = 0
    with torch.no_grad():
        for p in _trainable(model):
            n = p.numel()
            p.data.copy_(vec[off:off + n].view_as(p))
            off += n


def get_grad_vec(model) -> torch.Tensor:
    """utils.py:261-273 (missing grads read as zeros)."""
    fp = getattr(model, "_flat_params", None)
    if fp is not None:
        return fp.grad.clone()
    chunks = []
    for p in _trainable(model):
        if p.grad is None:
            chunks.append(torch.zeros(p.numel(), dtype=torch.float32, device=p.device))
        else:
            chunks.append(p.grad.detach().reshape(-1).float())
    return torch.cat(chunks)


def get_grad(model, args) -> torch.Tensor:
    """Gradient vector plus ``weight_decay / num_workers * w`` (utils.py:254-259;
    the division by the per-round client count is reference quirk Q7, kept
    because it changes the training curve)."""
    g = get_grad_vec(model)
    if args.weight_decay != 0:
        g.add_(get_param_vec(model), alpha=args.weight_decay / args.num_workers)
    return g.to(args.device)


def zero_grad(model) -> None:
    fp = getattr(model, "_flat_params", None)
    if fp is not None:
        fp.zero_grad()
        return
    for p in model.parameters():
        if p.grad is not None:
            p.grad.detach_()
            p.grad.zero_()


def clip_grad(l2_norm_clip: float, record):
    """Scale ``record`` (tensor or CSVec) down to L2 norm ``l2_norm_clip`` if it
    is larger (utils.py:305-313).  A CSVec is measured with ``l2estimate()``."""
    if torch.is_tensor(record):
        norm = float(torch.linalg.vector_norm(record))
    else:
        norm = float(record.l2estimate())
    if norm < l2_norm_clip:
        return record
    return record / abs(norm / l2_norm_clip)
