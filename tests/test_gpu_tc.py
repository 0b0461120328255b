"""tcgen05 / TMEM / TMA kernels vs fp32 PyTorch references of the same ops."""
import pytest
import torch
import torch.nn.functional as F

from commefficient_b200.ops import tc

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rel_err(got, want):
    return float((got.float() - want.float()).norm() / want.float().norm().clamp_min(1e-12))


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 128, 256), (128, 64, 128), (1000, 768, 768),
                                   (4096, 2304, 768), (77, 200, 72), (512, 512, 4608), (300, 50264, 64)])
def test_gemm_tn_matches_fp32(M, N, K):
    g = torch.Generator(device=DEV).manual_seed(M + N + K)
    a = torch.randn(M, K, device=DEV, generator=g).bfloat16()
    b = torch.randn(N, K, device=DEV, generator=g).bfloat16()
    want = a.float() @ b.float().t()
    got = tc.gemm_tn(a, b, out_dtype=torch.float32)
    assert _rel_err(got, want) < 1e-5
    got16 = tc.gemm_tn(a, b)
    assert _rel_err(got16, want) < 5e-3


def test_gemm_bias_relu_accumulate():
    a = torch.randn(256, 128, device=DEV).bfloat16()
    b = torch.randn(192, 128, device=DEV).bfloat16()
    bias = torch.randn(192, device=DEV)
    want = torch.relu(a.float() @ b.float().t() + bias)
    got = tc.gemm_tn(a, b, bias=bias, relu=True, out_dtype=torch.float32)
    assert _rel_err(got, want) < 1e-5
    acc = torch.ones(256, 192, device=DEV)
    tc.gemm_tn(a, b, out=acc, accumulate=True)
    assert _rel_err(acc, 1 + a.float() @ b.float().t()) < 1e-5


@pytest.mark.parametrize("n,h,cin,cout", [(4, 32, 64, 128), (2, 16, 128, 128), (6, 8, 256, 512), (16, 4, 512, 512),
                                          (5, 16, 128, 256), (3, 4, 64, 64), (500, 32, 64, 128)])
@pytest.mark.parametrize("relu", [False, True])
def test_conv3x3_matches_fp32(n, h, cin, cout, relu):
    g = torch.Generator(device=DEV).manual_seed(n * h + cin)
    x = torch.randn(n, h, h, cin, device=DEV, generator=g).bfloat16()
    w = (torch.randn(cout, 3, 3, cin, device=DEV, generator=g) / (3 * cin ** 0.5)).bfloat16()
    want = F.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), padding=1).permute(0, 2, 3, 1)
    if relu:
        want = torch.relu(want)
    got = tc.conv3x3_nhwc(x, w, relu=relu, out_dtype=torch.float32)
    assert got.shape == want.shape
    assert _rel_err(got, want) < 1e-4


@pytest.mark.parametrize("n,h,cin,cout", [(4, 32, 64, 128), (2, 16, 128, 128), (6, 8, 256, 512), (16, 4, 512, 512),
                                          (5, 16, 128, 256), (3, 4, 64, 64), (100, 32, 64, 128)])
def test_conv3x3_wgrad_matches_fp32(n, h, cin, cout):
    g = torch.Generator(device=DEV).manual_seed(n * h + cout)
    x = torch.randn(n, h, h, cin, device=DEV, generator=g).bfloat16()
    dz = torch.randn(n, h, h, cout, device=DEV, generator=g).bfloat16()
    want = torch.nn.grad.conv2d_weight(x.float().permute(0, 3, 1, 2), (cout, cin, 3, 3),
                                       dz.float().permute(0, 3, 1, 2), padding=1).permute(0, 2, 3, 1)
    dw = torch.zeros(cout, 3, 3, cin, device=DEV)
    tc.conv3x3_wgrad_nhwc(dz, x, dw)
    assert _rel_err(dw, want) < 1e-4
    tc.conv3x3_wgrad_nhwc(dz, x, dw)          # accumulates
    assert _rel_err(dw, 2 * want) < 1e-4


def test_pool_and_relu_bwd_kernels():
    y = torch.relu(torch.randn(6, 8, 8, 64, device=DEV)).bfloat16()
    for k in (2, 4):
        want = F.max_pool2d(y.float().permute(0, 3, 1, 2), k).permute(0, 2, 3, 1)
        got = tc.maxpool_nhwc(y, k)
        assert torch.equal(got.float(), want)
        dp = torch.randn_like(got)
        yy = y.float().permute(0, 3, 1, 2).requires_grad_(True)
        F.max_pool2d(yy, k).backward(dp.float().permute(0, 3, 1, 2))
        want_dz = yy.grad.permute(0, 2, 3, 1) * (y.float() > 0)
        got_dz = tc.maxpool_bwd_nhwc(y, dp, k, relu_mask=True)
        assert _rel_err(got_dz, want_dz) < 1e-6
    dy = torch.randn_like(y)
    assert torch.equal(tc.relu_bwd(y, dy).float(), dy.float() * (y.float() > 0))
    w = torch.randn(128, 3, 3, 64, device=DEV)
    wd = tc.dgrad_weights(w)
    assert torch.equal(wd.float(), w.flip(1, 2).permute(3, 1, 2, 0).bfloat16().float())


def test_resnet9_native_matches_library_path():
    from commefficient_b200 import models
    from commefficient_b200.ops import nn as nnops
    from commefficient_b200.utils import FlatParams
    torch.manual_seed(0)
    m = models.ResNet9(num_classes=10)
    fp = FlatParams(m, DEV, layout="nhwc")
    m._flat_params = fp
    x = torch.randn(24, 3, 32, 32, device=DEV).contiguous(memory_format=torch.channels_last)
    tgt = torch.randint(0, 10, (24,), device=DEV)

    def run(native_on):
        nnops.set_native(native_on)
        try:
            fp.zero_grad()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                logits = m(x).float()
            loss = F.cross_entropy(logits, tgt, reduction="sum")
            loss.backward()
            return logits.detach().clone(), fp.grad.clone()
        finally:
            nnops.set_native(False)

    l_lib, g_lib = run(False)
    l_nat, g_nat = run(True)
    assert _rel_err(l_nat, l_lib) < 3e-2
    assert _rel_err(g_nat, g_lib) < 0.1
    # per-layer check so a wrong layer cannot hide in the norm of the others
    for o, n_ in zip(fp.offsets, fp.sizes):
        assert _rel_err(g_nat[o:o + n_], g_lib[o:o + n_]) < 0.15, o


def test_mn_major_gemm_and_stem_im2col():
    a = torch.randn(1000, 128, device=DEV).bfloat16()       # [K, M]
    b = torch.randn(1000, 192, device=DEV).bfloat16()       # [K, N]
    c = torch.ones(128, 192, device=DEV)
    tc.gemm_mn_acc(a, b, c)
    assert _rel_err(c, 1 + a.float().t() @ b.float()) < 1e-5
    x = torch.randn(5, 32, 32, 3, device=DEV)
    col = tc.im2col_c3(x)
    want = F.unfold(x.permute(0, 3, 1, 2), 3, padding=1)     # [N, ci*9 (ci-major), L]
    want = want.view(5, 3, 9, 1024).permute(0, 3, 2, 1).reshape(5 * 1024, 27)   # -> column = tap*3 + ci
    assert torch.equal(col[:, :27].float(), want.bfloat16().float()) and float(col[:, 27:].abs().sum()) == 0


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (1000, 2304, 768), (300, 768, 3072), (77, 200, 72), (512, 64, 4608)])
def test_gemm_nn_matches_fp32(M, N, K):
    g = torch.Generator(device=DEV).manual_seed(M * 3 + N + K)
    a = torch.randn(M, K, device=DEV, generator=g).bfloat16()
    b = torch.randn(K, N, device=DEV, generator=g).bfloat16()
    bias = torch.randn(N, device=DEV, generator=g)
    got = tc.gemm_nn(a, b, bias=bias, out_dtype=torch.float32)
    assert _rel_err(got, a.float() @ b.float() + bias) < 1e-5


def test_gpt2_linears_native_match_library():
    from commefficient_b200 import models
    from commefficient_b200.ops import nn as nnops
    from commefficient_b200.utils import FlatParams
    torch.manual_seed(0)
    cfg = models.GPT2Config(vocab_size=512, n_positions=64, n_ctx=64, n_embd=128, n_layer=2, n_head=4,
                            resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0, summary_first_dropout=0.0)
    m = models.GPT2DoubleHeadsModel(cfg)
    fp = FlatParams(m, DEV)
    m._flat_params = fp
    ids = torch.randint(0, 512, (3, 2, 48), device=DEV)
    mc_tok = torch.full((3, 2), 47, device=DEV)
    mc_lab = torch.tensor([0, 1, 0], device=DEV)

    def run(native_on):
        nnops.set_native(native_on)
        try:
            fp.zero_grad()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                lm_loss, mc_loss, *_ = m(ids, token_type_ids=ids, mc_token_ids=mc_tok, lm_labels=ids, mc_labels=mc_lab)
            (lm_loss + mc_loss).backward()
            return float(lm_loss), float(mc_loss), fp.grad.clone()
        finally:
            nnops.set_native(False)

    l0, c0, g0 = run(False)
    l1, c1, g1 = run(True)
    assert abs(l0 - l1) < 2e-2 and abs(c0 - c1) < 2e-2
    assert _rel_err(g1, g0) < 0.08


def test_resnet18_stride1_convs_native_match_library():
    from commefficient_b200 import models
    from commefficient_b200.ops import nn as nnops
    from commefficient_b200.utils import FlatParams
    torch.manual_seed(0)
    m = models.FixupResNet18(num_classes=10)
    for p in m.parameters():                     # Fixup zero-inits half the convs; make every layer matter
        if p.dim() == 4 and float(p.abs().sum()) == 0:
            torch.nn.init.normal_(p, 0, 0.02)
    torch.nn.init.normal_(m.classifier.weight, 0, 0.05)
    fp = FlatParams(m, DEV, layout="nhwc")
    m._flat_params = fp
    x = torch.randn(8, 3, 32, 32, device=DEV).contiguous(memory_format=torch.channels_last)
    tgt = torch.randint(0, 10, (8,), device=DEV)

    def run(native_on):
        nnops.set_native(native_on)
        try:
            fp.zero_grad()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = F.cross_entropy(m(x).float(), tgt)
            loss.backward()
            return float(loss), fp.grad.clone()
        finally:
            nnops.set_native(False)

    l0, g0 = run(False)
    l1, g1 = run(True)
    assert abs(l0 - l1) < 2e-2
    assert _rel_err(g1, g0) < 0.08


def test_ce_layernorm_gelu_kernels_match_torch():
    from commefficient_b200.ops import tfm
    g = torch.Generator(device=DEV).manual_seed(5)
    logits = (torch.randn(37, 50262, device=DEV, generator=g) * 3).bfloat16().requires_grad_(True)
    labels = torch.randint(0, 50262, (37,), device=DEV, generator=g)
    labels[::5] = -1
    rows = tfm.cross_entropy_rows(logits, labels, -1)
    want = F.cross_entropy(logits.detach().float().requires_grad_(True), labels, ignore_index=-1, reduction="none")
    torch.testing.assert_close(rows, want, rtol=1e-4, atol=1e-4)
    w = torch.randn(37, device=DEV)
    (rows * w).sum().backward()
    ref_in = logits.detach().float().requires_grad_(True)
    (F.cross_entropy(ref_in, labels, ignore_index=-1, reduction="none") * w).sum().backward()
    assert _rel_err(logits.grad, ref_in.grad) < 1e-2

    x = torch.randn(300, 768, device=DEV, requires_grad=True)
    ln = torch.nn.LayerNorm(768).to(DEV)
    torch.nn.init.normal_(ln.weight, 1.0, 0.1)
    torch.nn.init.normal_(ln.bias, 0.0, 0.1)
    y = tfm.layer_norm(x, ln.weight, ln.bias, ln.eps)
    yr = ln(x.detach())
    assert _rel_err(y, yr) < 5e-3
    dy = torch.randn(300, 768, device=DEV).bfloat16()
    y.backward(dy)
    xr = x.detach().clone().requires_grad_(True)
    ln2 = torch.nn.LayerNorm(768).to(DEV)
    ln2.load_state_dict(ln.state_dict())
    ln2(xr).backward(dy.float())
    assert _rel_err(x.grad, xr.grad) < 1e-3
    assert _rel_err(ln.weight.grad, ln2.weight.grad) < 1e-3 and _rel_err(ln.bias.grad, ln2.bias.grad) < 1e-3

    h = torch.randn(64, 3072, device=DEV).bfloat16().requires_grad_(True)
    out = tfm.gelu_tanh(h)
    assert _rel_err(out, F.gelu(h.detach().float(), approximate="tanh")) < 5e-3
    out.backward(torch.ones_like(out))
    hr = h.detach().float().requires_grad_(True)
    F.gelu(hr, approximate="tanh").sum().backward()
    assert _rel_err(h.grad, hr.grad) < 1e-2


@pytest.mark.parametrize("M,N,K", [(300, 203, 128), (129, 50262, 64), (64, 70, 200)])
def test_gemm_tn_row_padded_operands(M, N, K):
    """Arbitrary N / K: operands and result live in buffers whose rows are padded to 16 bytes."""
    g = torch.Generator(device=DEV).manual_seed(N)
    a = tc.padded_empty(M, K, torch.bfloat16, DEV)
    b = tc.padded_empty(N, K, torch.bfloat16, DEV)
    a.copy_(torch.randn(M, K, device=DEV, generator=g))
    b.copy_(torch.randn(N, K, device=DEV, generator=g))
    want = a.float() @ b.float().t()
    got = tc.gemm_tn(a, b)
    assert got.shape == (M, N) and got.stride(0) % 8 == 0
    assert _rel_err(got, want) < 5e-3
    got32 = tc.gemm_tn(a, b, out_dtype=torch.float32)
    assert _rel_err(got32, want) < 1e-5


@pytest.mark.parametrize("K,M,N", [(500, 203, 128), (300, 50262, 64), (64, 1, 64)])
def test_gemm_mn_acc_arbitrary_m(K, M, N):
    g = torch.Generator(device=DEV).manual_seed(M)
    a = tc.padded_empty(K, M, torch.bfloat16, DEV)
    a.copy_(torch.randn(K, M, device=DEV, generator=g))
    b = torch.randn(K, N, device=DEV, generator=g).bfloat16()
    c = torch.ones(M, N, device=DEV)
    tc.gemm_mn_acc(a, b, c)
    assert _rel_err(c, 1 + a.float().t() @ b.float()) < 1e-5


@pytest.mark.parametrize("R,C", [(768, 2304), (50262, 768), (65, 130), (3, 7)])
def test_cast_transpose(R, C):
    src = torch.randn(R, C, device=DEV)
    out = tc.cast_transpose_bf16(src)
    assert out.shape == (C, R) and out.stride(0) % 8 == 0
    assert torch.equal(out, src.t().bfloat16())


@pytest.mark.parametrize("rows,cols", [(16384, 768), (1000, 2304), (7, 8), (333, 3072)])
def test_colsum(rows, cols):
    x = torch.randn(rows, cols, device=DEV).bfloat16()
    out = torch.ones(cols, device=DEV)
    tc.colsum_acc(x, out)
    torch.testing.assert_close(out, 1 + x.float().sum(0), rtol=1e-4, atol=1e-3)


@pytest.mark.parametrize("n_out,conv1d", [(203, False), (50262, False), (192, True), (128, False)])
def test_linear_autograd_any_width(n_out, conv1d):
    """_LinearFn vs fp32 torch for nn.Linear and Conv1D storage, including widths that are not a
    multiple of 8 (row-padded outputs and gradients, the LM-head case)."""
    torch.manual_seed(n_out)
    K, M = 128, 96
    store = (torch.randn(K, n_out, device=DEV) * 0.05) if conv1d else (torch.randn(n_out, K, device=DEV) * 0.05)
    store.requires_grad_(True)
    weight = store.t() if conv1d else store
    bias = torch.randn(n_out, device=DEV, requires_grad=True) if n_out % 8 == 0 else None
    x = torch.randn(M, K, device=DEV).bfloat16().requires_grad_(True)
    assert tc.linear_supported(x, weight)
    y = tc.linear(x, weight, bias)
    assert y.shape == (M, n_out)
    dy = torch.randn(M, n_out, device=DEV)
    (y.float() * dy).sum().backward()
    xr = x.detach().float().requires_grad_(True)
    sr = store.detach().clone().requires_grad_(True)
    wr = sr.t() if conv1d else sr
    br = bias.detach().clone().requires_grad_(True) if bias is not None else None
    yr = F.linear(xr, wr, br)
    (yr * dy).sum().backward()
    assert _rel_err(y, yr) < 1e-2
    assert _rel_err(x.grad, xr.grad) < 2e-2
    assert _rel_err(store.grad, sr.grad) < 2e-2
    if bias is not None:
        assert _rel_err(bias.grad, br.grad) < 2e-2


def test_layernorm_bwd_widths():
    from commefficient_b200.ops import tfm
    for E in (128, 768, 1024, 96):
        x = torch.randn(517, E, device=DEV, requires_grad=True)
        ln = torch.nn.LayerNorm(E).to(DEV)
        torch.nn.init.normal_(ln.weight, 1.0, 0.1)
        y = tfm.layer_norm(x, ln.weight, ln.bias, ln.eps)
        dy = torch.randn(517, E, device=DEV).bfloat16()
        y.backward(dy)
        xr = x.detach().clone().requires_grad_(True)
        ln2 = torch.nn.LayerNorm(E).to(DEV)
        ln2.load_state_dict(ln.state_dict())
        ln2(xr).backward(dy.float())
        assert _rel_err(x.grad, xr.grad) < 1e-3, E
        assert _rel_err(ln.weight.grad, ln2.weight.grad) < 1e-3 and _rel_err(ln.bias.grad, ln2.bias.grad) < 1e-3, E


@pytest.mark.parametrize("R,V,E,frac", [(300, 50262, 128, 0.25), (513, 1000, 64, 0.6), (256, 777, 64, 0.0),
                                        (128, 512, 64, 1.0)])
def test_lm_head_ce_only_labeled_rows(R, V, E, frac):
    """Fused LM head + CE that skips unlabeled rows vs the dense fp32 computation: same per-row loss,
    same hidden / weight gradients."""
    g = torch.Generator(device=DEV).manual_seed(R + V)
    hidden = (torch.randn(R, E, device=DEV, generator=g) * 0.5).bfloat16().requires_grad_(True)
    weight = torch.nn.Parameter(torch.randn(V, E, device=DEV, generator=g) * 0.1)
    labels = torch.randint(0, V, (R,), device=DEV, generator=g)
    drop = torch.rand(R, device=DEV, generator=g) >= frac
    labels[drop] = -1
    wrow = torch.randn(R, device=DEV, generator=g)
    loss = tc.lm_head_ce(hidden, weight, labels, -1)
    (loss * wrow).sum().backward()

    hr = hidden.detach().float().requires_grad_(True)
    wr = weight.detach().bfloat16().float().requires_grad_(True)
    want = F.cross_entropy(hr @ wr.t(), labels, ignore_index=-1, reduction="none")
    (want * wrow).sum().backward()
    torch.testing.assert_close(loss, want, rtol=2e-2, atol=2e-2)
    assert bool((loss[labels == -1] == 0).all())
    if frac > 0:
        assert _rel_err(hidden.grad, hr.grad) < 3e-2
        assert _rel_err(weight.grad, wr.grad) < 3e-2
    else:
        assert float(hidden.grad.abs().max()) == 0 and float(weight.grad.abs().max()) == 0


def test_prepare_conv_weights_one_launch():
    ws = [torch.randn(co, 3, 3, ci, device=DEV) for co, ci in ((128, 64), (64, 64), (256, 128), (72, 8))]
    before = __import__("commefficient_b200.ops.native", fromlist=["x"]).launch_count()
    casts, rots = tc.prepare_conv_weights(ws)
    assert __import__("commefficient_b200.ops.native", fromlist=["x"]).launch_count() == before + 1
    for w, c, r in zip(ws, casts, rots):
        assert torch.equal(c, w.bfloat16())
        assert torch.equal(r, tc.dgrad_weights(w))


def test_linear_ce_head_kernel_matches_torch():
    g = torch.Generator(device=DEV).manual_seed(11)
    for N, F_, C in ((500, 512, 10), (37, 64, 32), (9, 1024, 3)):
        feat = torch.randn(N, F_, device=DEV, generator=g).bfloat16()
        W = torch.randn(C, F_, device=DEV, generator=g) * 0.2
        labels = torch.randint(0, C, (N,), device=DEV, generator=g)
        gvec = torch.rand(N, device=DEV, generator=g) + 0.5
        loss, correct = tc.linear_ce_head(feat, W, 0.125, labels)
        fr = feat.float().requires_grad_(True)
        Wr = W.clone().requires_grad_(True)
        logits = (fr @ Wr.t()) * 0.125
        want = F.cross_entropy(logits, labels, reduction="none")
        torch.testing.assert_close(loss, want, rtol=1e-4, atol=1e-4)
        assert torch.equal(correct, (logits.argmax(1) == labels).float())
        (want * gvec).sum().backward()
        dW = torch.ones(C, F_, device=DEV)
        dfeat = tc.linear_ce_head(feat, W, 0.125, labels, g=gvec, dW=dW)
        assert _rel_err(dfeat, fr.grad) < 1e-2
        assert _rel_err(dW - 1, Wr.grad) < 1e-4


def test_resnet9_fused_loss_matches_library_path():
    """ce_per_sample on the native path (trunk + fused head/CE kernel) vs the library path."""
    from commefficient_b200 import models
    from commefficient_b200.ops import nn as nnops
    from commefficient_b200.utils import FlatParams
    torch.manual_seed(1)
    m = models.ResNet9(num_classes=10)
    fp = FlatParams(m, DEV, layout="nhwc")
    m._flat_params = fp
    x = torch.randn(24, 3, 32, 32, device=DEV).contiguous(memory_format=torch.channels_last)
    tgt = torch.randint(0, 10, (24,), device=DEV)

    def run(native_on):
        nnops.set_native(native_on)
        try:
            fp.zero_grad()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss, correct = m.ce_per_sample(x, tgt)
            loss.float().sum().backward()
            return loss.detach().clone(), correct.detach().clone(), fp.grad.clone()
        finally:
            nnops.set_native(False)

    l_lib, c_lib, g_lib = run(False)
    l_nat, c_nat, g_nat = run(True)
    assert _rel_err(l_nat, l_lib) < 3e-2
    assert float((c_nat != c_lib).float().mean()) < 0.1      # near-ties may flip under bf16
    assert _rel_err(g_nat, g_lib) < 0.1
    for o, n_ in zip(fp.offsets, fp.sizes):
        assert _rel_err(g_nat[o:o + n_], g_lib[o:o + n_]) < 0.15, o


@pytest.mark.parametrize("n,h,cin,cout", [(5, 16, 128, 128), (7, 4, 512, 512), (3, 32, 64, 64)])
def test_conv3x3_fused_residual(n, h, cin, cout):
    g = torch.Generator(device=DEV).manual_seed(h + cin)
    x = torch.randn(n, h, h, cin, device=DEV, generator=g).bfloat16()
    w = (torch.randn(cout, 3, 3, cin, device=DEV, generator=g) / (3 * cin ** 0.5)).bfloat16()
    res = torch.randn(n, h, h, cout, device=DEV, generator=g).bfloat16()
    plain = tc.conv3x3_nhwc(x, w, relu=True)
    pre, post = tc.conv3x3_nhwc_res(x, w, res, relu=True, keep_pre_add=True)
    assert torch.equal(pre, plain) and torch.equal(post, plain + res)
    only = tc.conv3x3_nhwc_res(x, w, res, relu=False)
    assert torch.equal(only, tc.conv3x3_nhwc(x, w) + res)
    mask = torch.randn(n, h, h, cout, device=DEV, generator=g).bfloat16()
    masked = tc.conv3x3_nhwc_res(x, w, res, relu_mask=mask)
    assert torch.equal(masked, tc.conv3x3_nhwc(x, w, relu_mask=mask) + res)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("sizes,C,H", [([8] * 16, 64, 32), ([3, 5, 1, 7], 128, 8), ([2, 2, 2], 72, 5), ([6] * 4, 256, 4)])
def test_segment_batchnorm_native_matches_reference(sizes, C, H, dtype):
    """ceb_segment_bn(_f32)_fwd / bwd vs per-client nn.BatchNorm2d in fp32."""
    from commefficient_b200.ops import tfm
    torch.manual_seed(C + H)
    N, S = sum(sizes), len(sizes)
    lo = dtype == torch.bfloat16
    x = (torch.randn(N, C, H, H, device=DEV) * 1.5 + 0.3).to(dtype).contiguous(memory_format=torch.channels_last)
    x.requires_grad_(True)
    w = torch.nn.Parameter(torch.randn(C, device=DEV) * 0.2 + 1.0)
    b = torch.nn.Parameter(torch.randn(C, device=DEV) * 0.2)
    seg = torch.repeat_interleave(torch.arange(S, device=DEV), torch.tensor(sizes, device=DEV))
    samples = torch.tensor(sizes, device=DEV, dtype=torch.float32)
    assert tfm.segment_bn_supported(x, seg, samples)
    y, mean, var = tfm.segment_batch_norm(x, w, b, seg, samples, S, 1e-5)
    assert y.is_contiguous(memory_format=torch.channels_last)
    assert y.dtype == dtype
    dy = torch.randn(N, C, H, H, device=DEV).to(dtype).contiguous(memory_format=torch.channels_last)
    y.backward(dy)

    xr = x.detach().float().requires_grad_(True)
    ref = torch.nn.BatchNorm2d(C).to(DEV)
    with torch.no_grad():
        ref.weight.copy_(w)
        ref.bias.copy_(b)
    bounds = [0]
    for n_ in sizes:
        bounds.append(bounds[-1] + n_)
    yr = torch.cat([ref(xr[a:c]) for a, c in zip(bounds[:-1], bounds[1:])])
    yr.backward(dy.float())
    assert _rel_err(y, yr) < (1e-2 if lo else 1e-5)
    mr = torch.stack([xr[a:c].mean(dim=(0, 2, 3)) for a, c in zip(bounds[:-1], bounds[1:])])
    torch.testing.assert_close(mean, mr.detach(), rtol=1e-3, atol=1e-3)
    assert _rel_err(x.grad, xr.grad) < (2e-2 if lo else 1e-4)
    tol = 1e-2 if lo else 1e-4
    assert _rel_err(w.grad, ref.weight.grad) < tol and _rel_err(b.grad, ref.bias.grad) < tol


@pytest.mark.parametrize("E", [128, 768, 1024])
def test_add_layer_norm_fused_matches_torch(E):
    from commefficient_b200.ops import tfm
    torch.manual_seed(E)
    rows = 333
    resid = torch.randn(rows, E, device=DEV, requires_grad=True)
    branch = torch.randn(rows, E, device=DEV).bfloat16().requires_grad_(True)
    ln = torch.nn.LayerNorm(E).to(DEV)
    torch.nn.init.normal_(ln.weight, 1.0, 0.1)
    torch.nn.init.normal_(ln.bias, 0.0, 0.1)
    new_resid, y = tfm.add_layer_norm(resid, branch, ln.weight, ln.bias, ln.eps)
    d_res = torch.randn(rows, E, device=DEV)
    dy = torch.randn(rows, E, device=DEV).bfloat16()
    torch.autograd.backward([new_resid, y], [d_res, dy])

    r2 = resid.detach().clone().requires_grad_(True)
    b2 = branch.detach().float().requires_grad_(True)
    ln2 = torch.nn.LayerNorm(E).to(DEV)
    ln2.load_state_dict(ln.state_dict())
    nr = r2 + b2
    y2 = ln2(nr)
    torch.autograd.backward([nr, y2], [d_res, dy.float()])
    torch.testing.assert_close(new_resid, nr.detach(), rtol=1e-6, atol=1e-6)
    assert _rel_err(y, y2) < 5e-3
    assert _rel_err(resid.grad, r2.grad) < 1e-3
    assert _rel_err(branch.grad, b2.grad) < 5e-3
    assert _rel_err(ln.weight.grad, ln2.weight.grad) < 1e-3 and _rel_err(ln.bias.grad, ln2.bias.grad) < 1e-3
    # no gradient through the residual output (the last add of the network)
    resid.grad = None
    nr3, y3 = tfm.add_layer_norm(resid, branch, ln.weight, ln.bias, ln.eps)
    y3.backward(dy)
    r2.grad = None
    ln2(r2 + b2.detach()).backward(dy.float())
    assert _rel_err(resid.grad, r2.grad) < 1e-3


@pytest.mark.parametrize("M,N,K", [(4096, 768, 50264), (300, 256, 8000), (128, 64, 4096), (1000, 768, 768)])
def test_gemm_accumulate_split_k_matches_fp32(M, N, K):
    """fp32-accumulate outputs may split the reduction over CTAs (few output tiles, long K: the LM-head data
    gradient); the result is the plain product added to what was there."""
    g = torch.Generator(device=DEV).manual_seed(M + K)
    a = (torch.randn(M, K, device=DEV, generator=g) * 0.1).bfloat16()
    b = (torch.randn(N, K, device=DEV, generator=g) * 0.1).bfloat16()
    acc = torch.ones(M, N, device=DEV)
    tc.gemm_tn(a, b, out=acc, accumulate=True)
    assert _rel_err(acc, 1 + a.float() @ b.float().t()) < 2e-5


@pytest.mark.parametrize("M,K,N", [(1000, 768, 3072), (257, 128, 256), (4096, 256, 64)])
def test_gemm_gelu_epilogues_match_separate_kernels(M, K, N):
    """ceb_gemm_bf16_tn_gelu: act 2 == GEMM followed by the GELU kernel (bit-exact: the activation is computed
    from the rounded pre-activation); act 3 == GEMM then GELU backward within bf16 rounding; both vs fp32 torch."""
    from commefficient_b200.ops import tfm
    torch.manual_seed(M + N)
    a = (torch.randn(M, K, device=DEV) * 0.5).bfloat16()
    b = (torch.randn(N, K, device=DEV) * 0.05).bfloat16()
    bias = torch.randn(N, device=DEV) * 0.1
    pre, act = tc.gemm_tn_gelu(a, b, bias)
    pre_ref = tc.gemm_tn(a, b, bias=bias)
    assert torch.equal(pre, pre_ref)
    assert torch.equal(act, tfm.gelu_tanh(pre_ref))
    want = torch.nn.functional.gelu(a.float() @ b.float().t() + bias, approximate="tanh")
    assert _rel_err(act, want) < 1e-2

    g = (torch.randn(M, K, device=DEV) * 0.5).bfloat16()
    d = tc.gemm_tn_gelu_grad(g, b, pre)
    x = pre.float().requires_grad_(True)
    torch.nn.functional.gelu(x, approximate="tanh").backward(g.float() @ b.float().t())
    assert _rel_err(d, x.grad) < 1e-2


def test_mlp_gelu_fused_matches_unfused():
    """_MLPGeluFn (GELU inside the GEMM epilogues) against the linear -> gelu -> linear chain on the same kernels and
    against fp32 torch: outputs and all five gradients."""
    from commefficient_b200.ops import nn as nnops
    torch.manual_seed(5)
    E, rows = 256, 640
    x = torch.randn(rows, E, device=DEV)
    w_fc = torch.nn.Parameter(torch.randn(E, 4 * E, device=DEV) * 0.05)      # Conv1D storage [in, out]
    b_fc = torch.nn.Parameter(torch.randn(4 * E, device=DEV) * 0.1)
    w_pr = torch.nn.Parameter(torch.randn(4 * E, E, device=DEV) * 0.05)
    b_pr = torch.nn.Parameter(torch.randn(E, device=DEV) * 0.1)
    dy = torch.randn(rows, E, device=DEV)
    params = (w_fc, b_fc, w_pr, b_pr)

    def run(fn):
        for p in params:
            p.grad = None
        xi = x.clone().requires_grad_(True)
        y = fn(xi)
        y.float().backward(dy)
        return [y.float(), xi.grad.float()] + [p.grad.float().clone() for p in params]

    nnops.set_native(True)
    nnops.set_precision("bf16")
    assert tc.mlp_gelu_supported(x, w_fc.t(), w_pr.t())
    fused = run(lambda xi: nnops.mlp_gelu(xi, w_fc.t(), b_fc, w_pr.t(), b_pr))
    chain = run(lambda xi: nnops.linear(nnops.gelu_tanh(nnops.linear(xi, w_fc.t(), b_fc)), w_pr.t(), b_pr))
    ref = run(lambda xi: torch.nn.functional.gelu(xi @ w_fc + b_fc, approximate="tanh") @ w_pr + b_pr)
    for f, c, r in zip(fused, chain, ref):
        assert _rel_err(f, c) < 5e-3
        assert _rel_err(f, r) < 2e-2
