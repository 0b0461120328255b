"""Out-of-bounds canaries: every writing kernel gets its output placed between two guard
regions filled with a sentinel; after the call the guards must be untouched and the payload
fully written.  (compute-sanitizer is not available on the GPU pool, so this is the bounds
check the kernels are held to; odd sizes exercise every tail path.)"""
import pytest
import torch

from commefficient_b200.ops import native, tc, hashing
from commefficient_b200.ops import sketch as sk

pytestmark = pytest.mark.gpu
DEV = "cuda"
GUARD = 4096            # elements on each side
SENT = -7777.0


class Guarded:
    """A tensor view with sentinel-filled guard bands on both sides."""

    def __init__(self, numel, dtype):
        self.buf = torch.full((numel + 2 * GUARD,), SENT, dtype=dtype, device=DEV)
        self.numel = numel
        self.view = self.buf[GUARD:GUARD + numel]
        # payload starts as NaN so "not written" is visible too
        self.view.fill_(float("nan"))

    def check(self, written=True):
        torch.cuda.synchronize()
        lo, hi = self.buf[:GUARD], self.buf[GUARD + self.numel:]
        assert bool((lo == SENT).all()), "write before the start of the output"
        assert bool((hi == SENT).all()), "write past the end of the output"
        if written:
            assert not bool(torch.isnan(self.view.float()).any()), "output not fully written"


@pytest.mark.parametrize("M,N,K", [(77, 200, 72), (129, 72, 64), (1, 8, 8), (300, 1000, 136)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_gemm_tn_stays_in_bounds(M, N, K, dtype):
    a = torch.randn(M, K, device=DEV).bfloat16()
    b = torch.randn(N, K, device=DEV).bfloat16()
    g = Guarded(M * N, dtype)
    tc.gemm_tn(a, b, out=g.view.view(M, N))
    g.check()
    want = a.float() @ b.float().t()
    assert float((g.view.view(M, N).float() - want).norm() / want.norm()) < 1e-2


@pytest.mark.parametrize("n,h,cin,cout", [(3, 4, 64, 64), (5, 16, 128, 256), (7, 32, 64, 128), (1, 8, 256, 512)])
def test_conv_fprop_stays_in_bounds(n, h, cin, cout):
    x = torch.randn(n, h, h, cin, device=DEV).bfloat16()
    w = (torch.randn(cout, 3, 3, cin, device=DEV) / 30).bfloat16()
    g = Guarded(n * h * h * cout, torch.bfloat16)
    native.call("ceb_conv3x3_bf16_nhwc", native.ptr(x), native.ptr(w), native.ptr(g.view), n, h, h, cin, cout, None,
                1, 0, None, native.stream_ptr(x.device))
    g.check()


@pytest.mark.parametrize("n,h,cin,cout", [(3, 4, 64, 64), (5, 16, 128, 256), (2, 32, 64, 128)])
def test_conv_wgrad_stays_in_bounds(n, h, cin, cout):
    x = torch.randn(n, h, h, cin, device=DEV).bfloat16()
    dz = torch.randn(n, h, h, cout, device=DEV).bfloat16()
    g = Guarded(cout * 9 * cin, torch.float32)
    g.view.zero_()
    native.call("ceb_conv3x3_wgrad_bf16_nhwc", native.ptr(dz), native.ptr(x), native.ptr(g.view), n, h, h, cin, cout,
                native.stream_ptr(x.device))
    g.check()


@pytest.mark.parametrize("K,M,N", [(100, 64, 64), (333, 128, 192), (64, 64, 512), (7, 192, 64)])
def test_gemm_mn_acc_stays_in_bounds(K, M, N):
    a = torch.randn(K, M, device=DEV).bfloat16()
    b = torch.randn(K, N, device=DEV).bfloat16()
    g = Guarded(M * N, torch.float32)
    g.view.zero_()
    tc.gemm_mn_acc(a, b, g.view.view(M, N))
    g.check()
    want = a.float().t() @ b.float()
    assert float((g.view.view(M, N) - want).norm() / want.norm()) < 1e-4


@pytest.mark.parametrize("numel", [1, 7, 1023, 65537])
def test_cast_and_relu_bwd_stay_in_bounds(numel):
    src = torch.randn(numel, device=DEV)
    g = Guarded(numel, torch.bfloat16)
    native.call("ceb_cast_f32_bf16", native.ptr(src), native.ptr(g.view), numel, native.stream_ptr(src.device))
    g.check()
    n8 = numel // 8 * 8
    if n8 == 0:
        return
    y = torch.randn(n8, device=DEV).bfloat16()
    dy = torch.randn(n8, device=DEV).bfloat16()
    g2 = Guarded(n8, torch.bfloat16)
    native.call("ceb_relu_bwd_bf16", native.ptr(y), native.ptr(dy), native.ptr(g2.view), n8,
                native.stream_ptr(y.device))
    g2.check()
    assert torch.equal(g2.view, torch.where(y > 0, dy, torch.zeros_like(dy)))


def test_bad_shapes_are_rejected_not_run():
    a = torch.randn(64, 64, device=DEV).bfloat16()
    b = torch.randn(65, 64, device=DEV).bfloat16()
    c = torch.empty(64, 65, device=DEV, dtype=torch.bfloat16)
    with pytest.raises(native.NativeError):          # dense entry point: N must be a multiple of 8
        native.call("ceb_gemm_bf16_tn", native.ptr(a), native.ptr(b), native.ptr(c), 64, 65, 64, None, 0, 0,
                    native.stream_ptr(a.device))
    with pytest.raises(native.NativeError):          # strided entry point: ldc not a multiple of 8
        native.call("ceb_gemm_bf16_tn_ld", native.ptr(a), 64, native.ptr(b), 64, native.ptr(c), 65, 64, 65, 64, None,
                    0, 0, native.stream_ptr(a.device))
    with pytest.raises(native.NativeError):          # N of the MN-major GEMM must be a multiple of 64
        tc.gemm_mn_acc(torch.randn(8, 64, device=DEV).bfloat16(), torch.randn(8, 72, device=DEV).bfloat16(),
                       torch.zeros(64, 72, device=DEV))


@pytest.mark.parametrize("n,h,c,k", [(3, 32, 128, 2), (2, 8, 512, 2), (5, 4, 512, 4)])
def test_maxpool_stays_in_bounds(n, h, c, k):
    y = torch.randn(n, h, h, c, device=DEV).bfloat16()
    g = Guarded(n * (h // k) * (h // k) * c, torch.bfloat16)
    native.call("ceb_maxpool_nhwc_bf16", native.ptr(y), native.ptr(g.view), n, h, h, c, k,
                native.stream_ptr(y.device))
    g.check()
    dp = torch.randn(n, h // k, h // k, c, device=DEV).bfloat16()
    g2 = Guarded(y.numel(), torch.bfloat16)
    native.call("ceb_maxpool_bwd_nhwc_bf16", native.ptr(y), native.ptr(dp), native.ptr(g2.view), n, h, h, c, k, 1,
                native.stream_ptr(y.device))
    g2.check()


@pytest.mark.parametrize("d,c,r", [(1000, 37, 3), (65537, 1001, 5), (6_568_640, 500_000, 5)])
def test_sketch_accumulate_stays_in_bounds(d, c, r):
    plan = sk.SketchPlan(d, c, r, 20 if d > 100000 else 1, DEV)
    vec = torch.randn(d, device=DEV)
    g = Guarded(r * c, torch.float32)
    g.view.zero_()
    sk.accumulate(plan, g.view.view(r, c), vec)
    g.check()
    ref = torch.zeros(r, c, device=DEV)
    sk.accumulate(plan, ref, vec)
    assert torch.allclose(ref.flatten(), g.view, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("d,k", [(1000, 10), (65537, 5000), (1_000_003, 1), (6_568_640, 50_000)])
def test_topk_select_stays_in_bounds(d, k):
    vec = torch.randn(d, device=DEV)
    gi = Guarded(k, torch.float32)          # int32 payload viewed through a float32 guard buffer
    gv = Guarded(k, torch.float32)
    scratch = native.select_scratch(vec.device)
    native.call("ceb_topk_select", native.ptr(vec), d, k, native.ptr(scratch), scratch.numel(),
                native.ptr(gi.view), native.ptr(gv.view), native.stream_ptr(vec.device))
    gi.check(written=False)
    gv.check()
    sel = gi.view.view(torch.int32).long()
    assert int(sel.min()) >= 0 and int(sel.max()) < d and sel.unique().numel() == k
    thresh = vec.abs().topk(k).values.min()
    assert bool((vec.abs()[sel] >= thresh).all())
    assert torch.equal(vec[sel], gv.view)
