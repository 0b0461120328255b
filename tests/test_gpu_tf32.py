"""The TF32 tensor-core path (``--precision tf32``: fp32 activations, tcgen05.mma.kind::tf32) against fp32
PyTorch references.  Operands are pre-truncated to TF32's 10 mantissa bits where the test wants an exact
comparison (products of TF32 numbers accumulate exactly like the fp32 reference up to summation order);
un-truncated operands are checked against the TF32 rounding bound."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from commefficient_b200.ops import tc

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rel_err(got, want):
    return float((got.float() - want.float()).norm() / want.float().norm().clamp_min(1e-12))


def _tf32(t):
    """Truncate to TF32 (keeps sign, 8 exponent bits, 10 mantissa bits)."""
    return (t.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)


@pytest.fixture(autouse=True)
def _ieee_reference():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    tc.set_precision("tf32")
    yield
    tc.set_precision("bf16")
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


@pytest.mark.parametrize("M,N,K", [(128, 128, 32), (256, 128, 256), (128, 64, 128), (1000, 768, 768),
                                   (4096, 2304, 768), (77, 200, 72), (512, 512, 4608), (512000, 64, 32)])
def test_gemm_tn_tf32(M, N, K):
    g = torch.Generator(device=DEV).manual_seed(M + N + K)
    a = _tf32(torch.randn(M, K, device=DEV, generator=g))
    b = _tf32(torch.randn(N, K, device=DEV, generator=g))
    want = a @ b.t()
    got = tc.gemm_tn(a, b)
    # (the tensor cores' fp32 accumulator is not round-to-nearest: long reductions drift by a few 1e-6)
    assert got.dtype == torch.float32 and _rel_err(got, want) < 3e-5
    a2 = torch.randn(M, K, device=DEV, generator=g)
    assert _rel_err(tc.gemm_tn(a2, b), a2 @ b.t()) < 2e-3          # raw fp32 in: TF32 rounding only


def test_gemm_tf32_bias_relu_accumulate():
    a = _tf32(torch.randn(256, 128, device=DEV))
    b = _tf32(torch.randn(192, 128, device=DEV))
    bias = torch.randn(192, device=DEV)
    got = tc.gemm_tn(a, b, bias=bias, relu=True)
    assert _rel_err(got, torch.relu(a @ b.t() + bias)) < 1e-5
    acc = torch.ones(256, 192, device=DEV)
    tc.gemm_tn(a, b, out=acc, accumulate=True)
    assert _rel_err(acc, 1 + a @ b.t()) < 1e-5


@pytest.mark.parametrize("n,h,cin,cout", [(4, 32, 64, 128), (2, 16, 128, 128), (6, 8, 256, 512), (16, 4, 512, 512),
                                          (5, 16, 128, 256), (3, 4, 64, 64), (500, 32, 64, 128), (100, 16, 128, 128)])
@pytest.mark.parametrize("relu", [False, True])
def test_conv3x3_tf32(n, h, cin, cout, relu):
    g = torch.Generator(device=DEV).manual_seed(n * h + cin)
    x = _tf32(torch.randn(n, h, h, cin, device=DEV, generator=g))
    w = _tf32(torch.randn(cout, 3, 3, cin, device=DEV, generator=g) / (3 * cin ** 0.5))
    want = F.conv2d(x.permute(0, 3, 1, 2), w.permute(0, 3, 1, 2), padding=1).permute(0, 2, 3, 1)
    if relu:
        want = torch.relu(want)
    got = tc.conv3x3_nhwc(x, w, relu=relu)
    assert got.shape == want.shape and got.dtype == torch.float32
    assert _rel_err(got, want) < 1e-5


def test_conv3x3_tf32_mask_and_residual_epilogues():
    g = torch.Generator(device=DEV).manual_seed(5)
    x = _tf32(torch.randn(8, 16, 16, 128, device=DEV, generator=g))
    w = _tf32(torch.randn(128, 3, 3, 128, device=DEV, generator=g) / 30)
    mask = torch.randn(8, 16, 16, 128, device=DEV, generator=g)
    resid = torch.randn(8, 16, 16, 128, device=DEV, generator=g)
    conv = F.conv2d(x.permute(0, 3, 1, 2), w.permute(0, 3, 1, 2), padding=1).permute(0, 2, 3, 1)
    got = tc.conv3x3_nhwc(x, w, relu_mask=mask)
    assert _rel_err(got, conv * (mask > 0)) < 1e-5
    y, y2 = tc.conv3x3_nhwc_res(x, w, resid, relu=True, keep_pre_add=True)
    assert _rel_err(y, torch.relu(conv)) < 1e-5 and _rel_err(y2, torch.relu(conv) + resid) < 1e-5
    y3 = tc.conv3x3_nhwc_res(x, w, resid)
    assert _rel_err(y3, conv + resid) < 1e-5


@pytest.mark.parametrize("n,h,cin,cout", [(4, 32, 64, 128), (2, 16, 128, 128), (6, 8, 256, 512), (16, 4, 512, 512),
                                          (5, 16, 128, 256), (3, 4, 64, 64), (100, 32, 64, 128)])
def test_conv3x3_wgrad_tf32(n, h, cin, cout):
    g = torch.Generator(device=DEV).manual_seed(n * h + cout)
    x = _tf32(torch.randn(n, h, h, cin, device=DEV, generator=g))
    dz = _tf32(torch.randn(n, h, h, cout, device=DEV, generator=g))
    want = torch.nn.grad.conv2d_weight(x.permute(0, 3, 1, 2), (cout, cin, 3, 3), dz.permute(0, 3, 1, 2),
                                       padding=1).permute(0, 2, 3, 1)
    dw = torch.zeros(cout, 3, 3, cin, device=DEV)
    tc.conv3x3_wgrad_nhwc(dz, x, dw)
    assert _rel_err(dw, want) < 1e-5
    tc.conv3x3_wgrad_nhwc(dz, x, dw)          # accumulates
    assert _rel_err(dw, 2 * want) < 1e-5


@pytest.mark.parametrize("K,M,N", [(1000, 128, 192), (4096, 768, 2304), (333, 64, 32), (20000, 64, 32), (2048, 300, 256)])
def test_mn_major_gemm_tf32(K, M, N):
    g = torch.Generator(device=DEV).manual_seed(K + M + N)
    a = _tf32(torch.randn(K, M + (-M % 4), device=DEV, generator=g))[:, :M]      # [K, M], 16-byte rows
    b = _tf32(torch.randn(K, N, device=DEV, generator=g))
    c = torch.ones(M, N, device=DEV)
    tc.gemm_mn_acc(a, b, c)
    assert _rel_err(c, 1 + a.t() @ b) < 1e-5


def test_fp32_pool_relu_im2col_dgrad_layout_kernels():
    y = torch.relu(torch.randn(6, 8, 8, 64, device=DEV))
    for k in (2, 4):
        want = F.max_pool2d(y.permute(0, 3, 1, 2), k).permute(0, 2, 3, 1)
        got = tc.maxpool_nhwc(y, k)
        assert got.dtype == torch.float32 and torch.equal(got, want)
        dp = torch.randn_like(got)
        yy = y.permute(0, 3, 1, 2).clone().requires_grad_(True)
        F.max_pool2d(yy, k).backward(dp.permute(0, 3, 1, 2))
        assert torch.equal(tc.maxpool_bwd_nhwc(y, dp, k, relu_mask=True), yy.grad.permute(0, 2, 3, 1) * (y > 0))
    dy = torch.randn_like(y)
    assert torch.equal(tc.relu_bwd(y, dy), dy * (y > 0))
    w = torch.randn(128, 3, 3, 64, device=DEV)
    assert torch.equal(tc.dgrad_weights(w, dtype=torch.float32), w.flip(1, 2).permute(3, 1, 2, 0).contiguous())
    for cin in (1, 3):
        x = torch.randn(5, 28 if cin == 1 else 32, 32, cin, device=DEV)
        H, W = x.shape[1], x.shape[2]
        want = F.unfold(x.permute(0, 3, 1, 2), 3, padding=1).view(5, cin, 9, H * W).permute(0, 3, 2, 1)
        want = want.reshape(5 * H * W, 9 * cin)
        col = tc.im2col_c3(x, torch.float32)
        assert col.shape == (5 * H * W, 32) and torch.equal(col[:, :9 * cin], want)
        assert float(col[:, 9 * cin:].abs().sum()) == 0
        col16 = tc.im2col_c3(x, torch.bfloat16)
        assert torch.equal(col16[:, :9 * cin].float(), want.bfloat16().float())


def test_linear_ce_head_fp32_features():
    feat = torch.randn(300, 512, device=DEV)
    W = torch.randn(10, 512, device=DEV) * 0.05
    labels = torch.randint(0, 10, (300,), device=DEV)
    f2 = feat.clone().requires_grad_(True)
    W2 = W.clone().requires_grad_(True)
    logits = (f2 @ W2.t()) * 0.125
    want = F.cross_entropy(logits, labels, reduction="none")
    loss, correct = tc.linear_ce_head(feat, W, 0.125, labels)
    assert _rel_err(loss, want) < 1e-5
    assert torch.equal(correct, (logits.argmax(1) == labels).float())
    g = torch.rand(300, device=DEV)
    (want * g).sum().backward()
    dW = torch.zeros_like(W)
    df = tc.linear_ce_head(feat, W, 0.125, labels, g=g, dW=dW)
    assert df.dtype == torch.float32 and _rel_err(df, f2.grad) < 1e-5 and _rel_err(dW, W2.grad) < 1e-4


def test_resnet9_native_tf32_is_as_accurate_as_cudnn_tf32():
    """Whole network (loss and every layer's gradient).  The yardstick is what the REFERENCE gets on this
    hardware: its fp32 model runs its convolutions through cuDNN with TF32 allowed (the PyTorch default).
    Against IEEE fp32 (TF32 off) the native kind::tf32 path must be no further away than that library
    TF32 path is -- i.e. same precision class, not merely "close"."""
    from commefficient_b200 import models
    from commefficient_b200.ops import nn as nnops
    from commefficient_b200.utils import FlatParams
    torch.manual_seed(0)
    m = models.ResNet9(num_classes=10)
    fp = FlatParams(m, DEV, layout="nhwc")
    m._flat_params = fp
    x = torch.randn(48, 3, 32, 32, device=DEV).contiguous(memory_format=torch.channels_last)
    tgt = torch.randint(0, 10, (48,), device=DEV)

    def run(native_on, precision):
        nnops.set_native(native_on)
        nnops.set_precision(precision)
        try:
            fp.zero_grad()
            loss, correct = m.ce_per_sample(x, tgt)
            loss.sum().backward()
            return loss.detach().clone(), fp.grad.clone()
        finally:
            nnops.set_native(False)
            nnops.set_precision("bf16")
            torch.backends.cudnn.allow_tf32 = False
            torch.backends.cuda.matmul.allow_tf32 = False

    l_ref, g_ref = run(False, "fp32")
    l_lib, g_lib = run(False, "tf32")
    l_nat, g_nat = run(True, "tf32")
    assert _rel_err(l_nat, l_ref) < 1e-3
    lib_err, nat_err = _rel_err(g_lib, g_ref), _rel_err(g_nat, g_ref)
    assert nat_err < max(1.5 * lib_err, 1e-3), (nat_err, lib_err)
    assert nat_err < 0.05
    for o, n_ in zip(fp.offsets, fp.sizes):            # per layer: a wrong layer cannot hide in the others' norm
        lib_l, nat_l = _rel_err(g_lib[o:o + n_], g_ref[o:o + n_]), _rel_err(g_nat[o:o + n_], g_ref[o:o + n_])
        assert nat_l < max(2.0 * lib_l, 5e-3), (o, nat_l, lib_l)


def test_federated_rounds_tf32_native_match_cpu_fp32():
    """End to end through FedModel / FedOptimizer: the FULL-WIDTH ResNet-9 (so every convolution runs on
    the tcgen05 TF32 kernels, CUDA graph from round 3) tracks the same rounds in fp32 on the CPU."""
    from commefficient_b200 import models
    from commefficient_b200.cv_train import compute_loss_train, compute_loss_val
    from commefficient_b200.data_utils import synthetic_image_batch
    from commefficient_b200.fed_aggregator import FedModel, FedOptimizer
    from commefficient_b200.ops import native
    from commefficient_b200.utils import make_args

    def run(device, precision):
        torch.manual_seed(0)
        args = make_args(mode="uncompressed", num_workers=4, num_clients=32, local_batch_size=-1, device=device,
                         dataset_name="CIFAR10", precision=precision, weight_decay=5e-4, local_momentum=0.0,
                         virtual_momentum=0.9, error_type="none")
        model = models.ResNet9(num_classes=10)
        fed = FedModel(model, compute_loss_train, args, compute_loss_val)
        opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args)
        opt.param_groups[0]["lr"] = 0.02
        losses = []
        for step in range(4):
            fed.train(True)
            out = fed(synthetic_image_batch(4, 4, seed=step, first_client=4 * step))
            opt.step()
            losses.append(float(np.mean(out[0])))
        w = fed.fp.to_reference_order(fed.fp.flat).detach().cpu()
        fed.rt.shutdown()
        return w, losses

    native.reset_launch_count()
    wg, lg = run("cuda", "tf32")
    assert native.launch_count() > 100, "the tensor-core kernels did not run"
    wc, lc = run("cpu", "fp32")
    np.testing.assert_allclose(lg, lc, rtol=3e-3, atol=3e-3)
    assert float((wg - wc).norm() / wc.norm()) < 2e-3
