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
