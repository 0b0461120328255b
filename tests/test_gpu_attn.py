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


def _keep_rows(seed, salt, rows, cols, p_drop):
    """Keep-mask [rows, cols] of the row / column dropout hash (csrc/common.cuh) with torch integer ops."""
    M32 = 0xFFFFFFFF

    def h(x):
        x = x & M32
        x = x ^ (x >> 16); x = (x * 0x7feb352d) & M32
        x = x ^ (x >> 15); x = (x * 0x846ca68b) & M32
        return x ^ (x >> 16)
    seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    r = torch.arange(rows, dtype=torch.int64, device=DEV).view(rows, 1)
    rowkey = h((seed & M32) ^ h((seed >> 32) ^ (salt & M32) ^ h(r)))
    c = torch.arange(cols, dtype=torch.int64, device=DEV).view(1, cols)
    th = max(min(int(p_drop * 4294967296.0), M32), 1)
    return h(rowkey + c) >= th


@pytest.mark.parametrize("p_drop", [0.0, 0.1])
@pytest.mark.parametrize("with_tt", [False, True])
def test_embedding_kernels(p_drop, with_tt):
    V, P, E, N, T = 1000, 64, 768, 6, 48
    g = torch.Generator(device=DEV).manual_seed(3)
    wte = torch.nn.Parameter(torch.randn(V, E, device=DEV, generator=g))
    wpe = torch.nn.Parameter(torch.randn(P, E, device=DEV, generator=g))
    wte.grad, wpe.grad = torch.zeros_like(wte), torch.zeros_like(wpe)
    ids = torch.randint(0, V, (N, T), device=DEV, generator=g)
    tt = torch.randint(0, 3, (N, T), device=DEV, generator=g) if with_tt else None
    state = tfm.DropoutState(DEV, seed=11)
    state.advance()
    x = tfm.embed(wte, wpe, ids, tt, T, p_drop, 9, state)
    assert x.shape == (N, T, E) and x.dtype == torch.float32
    keep = _keep_rows(int(state.seed.item()), 9, N * T, E, p_drop).float().view(N, T, E) / (1 - p_drop) if p_drop else 1.0
    w2 = wte.detach().clone().requires_grad_(True)
    p2 = wpe.detach().clone().requires_grad_(True)
    want = w2[ids] + p2[torch.arange(T, device=DEV)]
    if with_tt:
        want = want + w2[tt]
    want = want * keep
    torch.testing.assert_close(x, want, rtol=1e-6, atol=1e-6)
    dx = torch.randn(N, T, E, device=DEV, generator=g)
    x.backward(dx)
    want.backward(dx)
    torch.testing.assert_close(wte.grad, w2.grad, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(wpe.grad, p2.grad, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("p_drop", [0.0, 0.1])
def test_add_layernorm_with_fused_dropout(p_drop):
    rows, E = 300, 768
    g = torch.Generator(device=DEV).manual_seed(4)
    resid = torch.randn(rows, E, device=DEV, generator=g).requires_grad_(True)
    branch = torch.randn(rows, E, device=DEV, generator=g).bfloat16().requires_grad_(True)
    w = torch.randn(E, device=DEV, generator=g).requires_grad_(True)
    b = torch.randn(E, device=DEV, generator=g).requires_grad_(True)
    state = tfm.DropoutState(DEV, seed=2)
    state.advance()
    new_resid, y = tfm.add_layer_norm(resid, branch, w, b, 1e-5, p_drop, 21, state)
    keep = _keep_rows(int(state.seed.item()), 21, rows, E, p_drop).float() / (1 - p_drop) if p_drop else 1.0
    r2, b2 = resid.detach().clone().requires_grad_(True), branch.detach().float().requires_grad_(True)
    w2, bb2 = w.detach().clone().requires_grad_(True), b.detach().clone().requires_grad_(True)
    want_r = r2 + b2 * keep
    want_y = torch.nn.functional.layer_norm(want_r, (E,), w2, bb2, 1e-5)
    torch.testing.assert_close(new_resid, want_r, rtol=1e-5, atol=1e-5)
    assert _rel(y, want_y) < 5e-3
    dy = torch.randn(rows, E, device=DEV, generator=g).bfloat16()
    dr = torch.randn(rows, E, device=DEV, generator=g)
    torch.autograd.backward([new_resid, y], [dr, dy])
    torch.autograd.backward([want_r, want_y], [dr, dy.float()])
    assert _rel(resid.grad, r2.grad) < 1e-4
    assert _rel(branch.grad, b2.grad) < 5e-3
    assert _rel(w.grad, w2.grad) < 1e-3 and _rel(b.grad, bb2.grad) < 1e-3


def test_gpt2_native_trunk_matches_library_without_dropout_and_runs_with_it():
    """Whole model on the native path (embedding, attention, fused add + LayerNorm, GEMMs, LM head) against
    the library path with dropout off; with dropout on the native path must give a finite, different loss
    whose masks change from one forward to the next."""
    from commefficient_b200 import models
    from commefficient_b200.ops import nn as nnops
    from commefficient_b200.utils import FlatParams
    torch.manual_seed(0)

    def build(pd):
        cfg = models.GPT2Config(vocab_size=512, n_positions=256, n_ctx=256, n_embd=128, n_layer=2, n_head=2,
                                resid_pdrop=pd, embd_pdrop=pd, attn_pdrop=pd, summary_first_dropout=0.0)
        torch.manual_seed(0)
        m = models.GPT2DoubleHeadsModel(cfg)
        fp = FlatParams(m, DEV)
        m._flat_params = fp
        return m, fp

    ids = torch.randint(0, 512, (3, 2, 160), device=DEV)
    mc_tok = torch.full((3, 2), 159, device=DEV)
    mc_lab = torch.tensor([0, 1, 0], device=DEV)

    def run(m, fp, native_on):
        nnops.set_native(native_on)
        nnops.set_precision("bf16")
        try:
            fp.zero_grad()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                lm_loss, mc_loss, *_ = m(ids, token_type_ids=ids, mc_token_ids=mc_tok, lm_labels=ids, mc_labels=mc_lab)
            (lm_loss + mc_loss).backward()
            return float(lm_loss.detach()), float(mc_loss.detach()), fp.grad.clone()
        finally:
            nnops.set_native(False)

    m, fp = build(0.0)
    l0, c0, g0 = run(m, fp, False)
    l1, c1, g1 = run(m, fp, True)
    assert abs(l0 - l1) < 2e-2 and abs(c0 - c1) < 2e-2
    assert _rel(g1, g0) < 0.08
    m, fp = build(0.1)
    la, _, ga = run(m, fp, True)
    lb, _, gb = run(m, fp, True)
    assert torch.isfinite(ga).all() and la != lb and abs(la - l1) < 0.5
