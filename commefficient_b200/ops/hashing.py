"""Count-sketch hash material, bit-compatible with the public ``csvec`` package
the reference imports (/root/reference/CommEfficient/fed_aggregator.py:5,464-467;
semantics restated in SURVEY.md section 2.2 -- the package itself is not on this
box).

For row ``r`` and token ``t`` (``t = i mod n`` for coordinate ``i`` in hash block
``b = i // n``, ``n = ceil(d / numBlocks)``):

    sign_r(t)   = ((((h2*t + h3)*t + h4)*t + h5) mod P mod 2) * 2 - 1     (4-wise)
    bucket_r(t) = ((h0*t + h1) mod P) mod c                               (2-wise)
    sign(i)     = sign_r(t) * blockSign[b]
    bucket(i)   = (bucket_r(t) + blockOffset[b]) mod c

with ``P = 2^61 - 1``, the six ``h`` drawn per row from ``randint(0, P)`` under
torch seed 42, int64 wrap-around arithmetic and floor-mod.  All random material
is drawn on the CPU generator so CPU and GPU sketches agree (the upstream
package draws block signs/offsets on the sketch's device, so its CUDA and CPU
sketches differ from each other; we pick the CPU stream for both).

The device kernels consume a *packed token table*: one ``int32`` per (row,
token) holding ``bucket_r(t) | (sign_r(t) < 0) << 31`` -- 4 bytes instead of the
upstream 12 (int64 bucket + fp32 sign), ``r * n * 4`` bytes in total (6.6 MB for
ResNet-9 at the default 20 blocks, so it stays L2-resident).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Tuple

import torch

LARGEPRIME = 2 ** 61 - 1
_SEED = 42


@dataclass
class SketchHashes:
    d: int
    c: int
    r: int
    num_blocks: int
    n_tokens: int
    buckets: torch.Tensor        # (r, n) int64, CPU
    signs: torch.Tensor          # (r, n) float32 in {-1, +1}, CPU
    block_offsets: torch.Tensor  # (num_blocks,) int64, CPU (zeros when num_blocks == 1)
    block_signs: torch.Tensor    # (num_blocks,) float32, CPU (ones when num_blocks == 1)
    packed: torch.Tensor         # (r, n) int32, CPU: bucket | signbit << 31

    def coord_buckets(self, row: int) -> torch.Tensor:
        """(d,) int64 bucket of every coordinate for ``row`` (oracle use)."""
        n = self.n_tokens
        i = torch.arange(self.d, dtype=torch.int64)
        t, b = i % n, i // n
        return (self.buckets[row, t] + self.block_offsets[b]) % self.c

    def coord_signs(self, row: int) -> torch.Tensor:
        n = self.n_tokens
        i = torch.arange(self.d, dtype=torch.int64)
        t, b = i % n, i // n
        return self.signs[row, t] * self.block_signs[b]


_CACHE: Dict[Tuple[int, int, int, int], SketchHashes] = {}


def make_hashes(d: int, c: int, r: int, num_blocks: int = 1) -> SketchHashes:
    key = (int(d), int(c), int(r), int(num_blocks))
    hit = _CACHE.get(key)
    if hit is not None:
        return hit
    d, c, r, num_blocks = key
    if c >= 2 ** 31:
        raise ValueError("num_cols must be < 2^31 (packed int32 bucket)")
    saved = torch.random.get_rng_state()
    try:
        torch.random.manual_seed(_SEED)
        h = torch.randint(0, LARGEPRIME, (r, 6), dtype=torch.int64, device="cpu")
        if num_blocks > 1:
            n = (d + num_blocks - 1) // num_blocks
            block_signs = (torch.randint(0, 2, (num_blocks,), device="cpu") * 2 - 1).float()
            block_offsets = torch.randint(0, c, (num_blocks,), device="cpu").long()
        else:
            n = d
            block_signs = torch.ones(1)
            block_offsets = torch.zeros(1, dtype=torch.int64)
    finally:
        torch.random.set_rng_state(saved)

    t = torch.arange(n, dtype=torch.int64).view(1, n)
    poly = ((h[:, 2:3] * t + h[:, 3:4]) * t + h[:, 4:5]) * t + h[:, 5:6]
    signs = ((poly % LARGEPRIME % 2) * 2 - 1).float()
    buckets = ((h[:, 0:1] * t) + h[:, 1:2]) % LARGEPRIME % c
    packed = buckets.to(torch.int32)
    neg = signs < 0
    # set bit 31 for negative signs (int32 two's complement: OR with INT_MIN)
    packed = torch.where(neg, packed | torch.tensor(-2 ** 31, dtype=torch.int32), packed)
    out = SketchHashes(d, c, r, num_blocks, n, buckets, signs, block_offsets,
                       block_signs, packed.contiguous())
    _CACHE[key] = out
    return out
