"""tcgen05 causal attention (csrc/attn.cu) against a plain fp32 PyTorch implementation of the same op:
forward output and all three input gradients, with and without dropout (the oracle rebuilds the kernel's
counter-hash mask), one tile and several tiles, sequence lengths that do not fill a tile."""
import math

import pytest
import torch

from commefficient_b200.ops import tfm

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rel(a, b):
    return float((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12))


def _reference(qkv, B, T, H, keep=None, inv_keep=1.0):
    E = qkv.shape[-1] // 3
    q, k, v = qkv.float().view(B, T, 3, H, 64).permute(2, 0, 3, 1, 4)          # each [B, H, T, 64]
    s = (q @ k.transpose(-1, -2)) / math.sqrt(64)
    causal = torch.ones(T, T, dtype=torch.bool, device=qkv.device).tril()
    s = s.masked_fill(~causal, float("-inf"))
    a = torch.softmax(s, dim=-1)
    if keep is not None:
        a = a * keep.float() * inv_keep
    return (a @ v).permute(0, 2, 1, 3).reshape(B * T, E)


@pytest.mark.parametrize("B,T,H", [(2, 128, 12), (3, 64, 4), (1, 200, 2), (2, 256, 3), (1, 1024, 2), (5, 37, 1), (2, 300, 12)])
@pytest.mark.parametrize("p_drop", [0.0, 0.1])
def test_attention_matches_fp32(B, T, H, p_drop):
    g = torch.Generator(device=DEV).manual_seed(B * 1000 + T + H)
    E = H * 64
    qkv = (torch.randn(B * T, 3 * E, device=DEV, generator=g) * 0.7).bfloat16().requires_grad_(True)
    dout = torch.randn(B * T, E, device=DEV, generator=g).bfloat16()
    state = tfm.DropoutState(DEV, seed=7)
    state.advance()
    out = tfm.attention(qkv, B, T, H, p_drop=p_drop, salt=5, state=state)
    out.backward(dout)
    got_grad = qkv.grad.clone()

    keep = None
    if p_drop > 0:
        keep = tfm.attention_keep_mask(int(state.seed.item()), 5, B, H, T, p_drop, device=DEV)
        frac = float(keep.float().mean())
        assert abs(frac - (1 - p_drop)) < 0.02, frac
    ref_in = qkv.detach().float().requires_grad_(True)
    want = _reference(ref_in, B, T, H, keep, 1.0 / (1.0 - p_drop))
    want.backward(dout.float())
    assert out.dtype == torch.bfloat16 and out.shape == (B * T, E)
    assert _rel(out, want) < 1e-2                       # bf16 P / O rounding
    E3 = 3 * E
    for name, lo in (("dq", 0), ("dk", E), ("dv", 2 * E)):
        assert _rel(got_grad[:, lo:lo + E], ref_in.grad[:, lo:lo + E]) < 2e-2, name
    assert got_grad.shape == (B * T, E3)


def test_attention_dropout_masks_change_per_round_and_per_site():
    B, T, H = 2, 128, 2
    qkv = torch.randn(B * T, 3 * H * 64, device=DEV).bfloat16()
    state = tfm.DropoutState(DEV, seed=1)
    state.advance()
    a = tfm.attention(qkv, B, T, H, p_drop=0.5, salt=1, state=state)
    b = tfm.attention(qkv, B, T, H, p_drop=0.5, salt=1, state=state)
    c = tfm.attention(qkv, B, T, H, p_drop=0.5, salt=2, state=state)
    state.advance()
    d = tfm.attention(qkv, B, T, H, p_drop=0.5, salt=1, state=state)
    assert torch.equal(a, b) and not torch.equal(a, c) and not torch.equal(a, d)
    e = tfm.attention(qkv, B, T, H, p_drop=0.0)
    f = tfm.attention(qkv, B, T, H, p_drop=0.0)
    assert torch.equal(e, f)
