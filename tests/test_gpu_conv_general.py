"""Strided (3x3 / s2, 1x1 / s2), 1x1 and odd-sized-map (28, 14, 7, 3 wide) convolutions on the tcgen05 tile and
weight-gradient kernels -- the shapes of ResNet-18 / FixupResNet18 (CIFAR) and of the FEMNIST ResNet-9 -- against
fp32 PyTorch, for both operand precisions."""
import pytest
import torch
import torch.nn.functional as F

from commefficient_b200.ops import tc

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rel(a, b):
    return float((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12))


def _tf32(t):
    return (t.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)


def _prep(t, dtype):
    return t.bfloat16() if dtype == torch.bfloat16 else _tf32(t)


@pytest.fixture(autouse=True)
def _ieee_reference():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


SHAPES = [  # n, h, cin, cout, ks, stride
    (8, 32, 64, 128, 3, 2), (6, 16, 128, 256, 3, 2), (10, 8, 256, 256, 3, 2), (8, 32, 64, 128, 1, 2), (6, 16, 128, 256, 1, 2),
    (7, 8, 256, 256, 1, 1), (5, 28, 64, 128, 3, 1), (9, 14, 128, 128, 3, 1), (11, 14, 128, 256, 3, 1), (20, 7, 256, 512, 3, 1),
    (33, 3, 512, 512, 3, 1), (4, 28, 64, 64, 3, 2), (3, 7, 256, 256, 3, 2), (100, 28, 64, 128, 3, 1),
]


@pytest.mark.parametrize("n,h,cin,cout,ks,stride", SHAPES)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_general_conv_forward(n, h, cin, cout, ks, stride, dtype):
    g = torch.Generator(device=DEV).manual_seed(n * h + cin + ks + stride)
    pad = 1 if ks == 3 else 0
    x = _prep(torch.randn(n, h, h, cin, device=DEV, generator=g), dtype)
    w = _prep(torch.randn(cout, ks, ks, cin, device=DEV, generator=g) / (ks * cin ** 0.5), dtype)
    want = F.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), stride=stride, padding=pad).permute(0, 2, 3, 1)
    got = tc.conv2d_nhwc(x, w, ks, stride, pad)
    assert got.shape == want.shape and got.dtype == dtype
    assert _rel(got, want) < (5e-3 if dtype == torch.bfloat16 else 1e-5)
    got_r = tc.conv2d_nhwc(x, w, ks, stride, pad, relu=True)
    assert _rel(got_r, torch.relu(want)) < (5e-3 if dtype == torch.bfloat16 else 1e-5)


@pytest.mark.parametrize("n,h,cin,cout,ks,stride", [s for s in SHAPES if s[4] == 3])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_general_conv_wgrad(n, h, cin, cout, ks, stride, dtype):
    ho = (h + 2 - 3) // stride + 1
    if dtype == torch.float32 and ho > 32:
        pytest.skip("TF32 k-step holds 32 pixels")
    g = torch.Generator(device=DEV).manual_seed(n * h + cout + stride)
    x = _prep(torch.randn(n, h, h, cin, device=DEV, generator=g), dtype)
    dz = _prep(torch.randn(n, ho, ho, cout, device=DEV, generator=g), dtype)
    want = torch.nn.grad.conv2d_weight(x.float().permute(0, 3, 1, 2), (cout, cin, 3, 3), dz.float().permute(0, 3, 1, 2),
                                       stride=stride, padding=1).permute(0, 2, 3, 1)
    dw = torch.zeros(cout, 3, 3, cin, device=DEV)
    tc.conv3x3_wgrad_strided(dz, x, dw, stride)
    assert _rel(dw, want) < 2e-5
    tc.conv3x3_wgrad_strided(dz, x, dw, stride)
    assert _rel(dw, 2 * want) < 2e-5


@pytest.mark.parametrize("n,h,cin,cout,ks,stride", [(8, 32, 64, 128, 3, 2), (6, 16, 128, 256, 1, 2), (7, 8, 256, 256, 1, 1),
                                                    (5, 28, 64, 128, 3, 1), (9, 14, 128, 128, 3, 2), (12, 7, 256, 256, 3, 1)])
@pytest.mark.parametrize("precision", ["bf16", "tf32"])
def test_conv_autograd_function_matches_torch(n, h, cin, cout, ks, stride, precision):
    tc.set_precision(precision)
    try:
        g = torch.Generator(device=DEV).manual_seed(n + h + cin)
        pad = 1 if ks == 3 else 0
        x = torch.randn(n, cin, h, h, device=DEV, generator=g).contiguous(memory_format=torch.channels_last)
        w = (torch.randn(cout, cin, ks, ks, device=DEV, generator=g) / (ks * cin ** 0.5)).contiguous(memory_format=torch.channels_last)
        if precision == "tf32":
            x, w = _tf32(x.permute(0, 2, 3, 1)).permute(0, 3, 1, 2), _tf32(w.permute(0, 2, 3, 1)).permute(0, 3, 1, 2)
        else:
            x, w = x.bfloat16().float(), w.bfloat16().float()
        x1, w1 = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
        x2, w2 = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
        assert tc.conv_supported(x1, w1, stride, pad)
        # (no ReLU here: a pre-activation within rounding of zero would flip its mask in one implementation only
        # and the comparison of the gradients would then measure that, not the kernels)
        y1 = tc.conv2d(x1, w1, None, stride, pad, relu=False)
        y2 = F.conv2d(x2, w2, stride=stride, padding=pad)
        yr = tc.conv2d(x1.detach(), w1.detach(), None, stride, pad, relu=True)
        assert _rel(yr, torch.relu(y2)) < (2e-2 if precision == "bf16" else 2e-3)
        dy = torch.randn(y2.shape, device=DEV, generator=g)
        if precision == "tf32":
            dy = _tf32(dy)
        y1.float().backward(dy)
        y2.backward(dy)
        tol = 2e-2 if precision == "bf16" else 2e-3
        assert _rel(y1, y2) < tol and _rel(x1.grad, x2.grad) < tol and _rel(w1.grad, w2.grad) < tol
    finally:
        tc.set_precision("bf16")


def test_pool_floor_mode_and_window_three_and_stride_helpers():
    for dtype in (torch.bfloat16, torch.float32):
        y = torch.relu(torch.randn(5, 7, 7, 64, device=DEV)).to(dtype)
        for k in (2, 3):
            want = F.max_pool2d(y.float().permute(0, 3, 1, 2), k).permute(0, 2, 3, 1)
            got = tc.maxpool_nhwc(y, k)
            assert got.shape == want.shape and torch.equal(got.float(), want)
            dp = torch.randn_like(got)
            yy = y.float().permute(0, 3, 1, 2).clone().requires_grad_(True)
            F.max_pool2d(yy, k).backward(dp.float().permute(0, 3, 1, 2))
            got_dz = tc.maxpool_bwd_nhwc(y, dp, k, relu_mask=True)
            assert _rel(got_dz, yy.grad.permute(0, 2, 3, 1) * (y.float() > 0)) < 1e-6
        x = torch.randn(3, 7, 9, 64, device=DEV).to(dtype)
        assert torch.equal(tc.subsample2(x), x[:, ::2, ::2, :])
        gsmall = torch.randn(3, 4, 5, 64, device=DEV).to(dtype)
        up = tc.upsample2_zero(gsmall, 7, 9)
        want = torch.zeros(3, 7, 9, 64, device=DEV, dtype=dtype)
        want[:, ::2, ::2, :] = gsmall
        assert torch.equal(up, want)


@pytest.mark.parametrize("C", [62, 100])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_classifier_head_many_classes(C, dtype):
    feat = torch.randn(300, 512, device=DEV).to(dtype)
    W = torch.randn(C, 512, device=DEV) * 0.05
    labels = torch.randint(0, C, (300,), device=DEV)
    f2 = feat.float().clone().requires_grad_(True)
    W2 = W.clone().requires_grad_(True)
    logits = (f2 @ W2.t()) * 0.125
    want = F.cross_entropy(logits, labels, reduction="none")
    loss, correct = tc.linear_ce_head(feat, W, 0.125, labels)
    assert _rel(loss, want) < 1e-5
    assert torch.equal(correct, (logits.argmax(1) == labels).float())
    g = torch.rand(300, device=DEV)
    (want * g).sum().backward()
    dW = torch.zeros_like(W)
    df = tc.linear_ce_head(feat, W, 0.125, labels, g=g, dW=dW)
    assert _rel(df, f2.grad) < (1e-2 if dtype == torch.bfloat16 else 1e-5) and _rel(dW, W2.grad) < 1e-4


@pytest.mark.parametrize("precision", ["bf16", "tf32"])
def test_femnist_resnet9_native_matches_library(precision):
    """1 x 28 x 28 input, 62 classes: maps 28 / 14 / 7 / 3, im2col'd one-channel stem, floor-mode pooling, the
    62-class fused head -- every layer on the native kernels."""
    from commefficient_b200 import models
    from commefficient_b200.ops import native
    from commefficient_b200.ops import nn as nnops
    from commefficient_b200.utils import FlatParams
    torch.manual_seed(0)
    m = models.ResNet9(num_classes=62, initial_channels=1)
    fp = FlatParams(m, DEV, layout="nhwc")
    m._flat_params = fp
    x = torch.randn(40, 1, 28, 28, device=DEV).contiguous(memory_format=torch.channels_last)
    tgt = torch.randint(0, 62, (40,), device=DEV)

    def run(native_on, prec):
        nnops.set_native(native_on)
        nnops.set_precision(prec)
        try:
            fp.zero_grad()
            if prec == "bf16":
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    loss, correct = m.ce_per_sample(x, tgt)
            else:
                loss, correct = m.ce_per_sample(x, tgt)
            loss.float().sum().backward()
            return loss.detach().float().clone(), fp.grad.clone()
        finally:
            nnops.set_native(False)
            nnops.set_precision("bf16")
            torch.backends.cudnn.allow_tf32 = False
            torch.backends.cuda.matmul.allow_tf32 = False

    l_ref, g_ref = run(False, "fp32")
    l_lib, g_lib = run(False, precision)
    native.reset_launch_count()
    l_nat, g_nat = run(True, precision)
    assert native.launch_count() > 30, "the native network did not run"
    assert _rel(l_nat, l_ref) < (3e-2 if precision == "bf16" else 1e-3)
    lib_err, nat_err = _rel(g_lib, g_ref), _rel(g_nat, g_ref)
    assert nat_err < max(1.5 * lib_err, 2e-3), (nat_err, lib_err)
