"""ceb_sketch_unsketch_topk_static -- heavy-hitter recovery for callers that cannot
allocate device memory.

Who that is: a host program that embeds libceb200.so but does not own an allocator it may
call on the hot path -- its GPU memory was split up once at start-up, or it runs inside
another framework's memory pool, or it must never trigger the device-wide synchronisation
a cudaMalloc can cause.  The regular entry point asks such a caller for a d-float estimate
buffer and a select block; this variant keeps both in an array compiled into the library
(CEB_STATIC_UNSKETCH_MAX_D coordinates at most) and takes inputs and outputs only.

The test drives both entry points through the C ABI with ctypes (not through the Python op
layer) and requires identical selections on two consecutive calls with different sketches --
the second call sees the first call's leftovers in the shared array, so it also checks that
what must start at zero is cleared.  This file sorts last so that a run that stops at the
first failure has already executed every other test.
"""
import ctypes

import pytest
import torch

from commefficient_b200.ops import native
from commefficient_b200.ops import sketch as sk

pytestmark = pytest.mark.gpu

MAX_D = 8 * 1024 * 1024
c_i64, c_i32, c_ptr = ctypes.c_int64, ctypes.c_int, ctypes.c_void_p


def _entry_points():
    lib = native.lib()
    static = lib.ceb_sketch_unsketch_topk_static
    static.argtypes = [c_ptr, c_ptr, c_i64, c_ptr, c_i32, c_i32, c_i64, c_i64, c_i64, c_ptr, c_ptr, c_ptr]
    static.restype = c_i32
    return lib.ceb_sketch_unsketch_topk, static


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _sorted_by_index(idx, val):
    order = torch.argsort(idx)
    return idx[order], val[order]


def test_static_variant_matches_regular_entry_point_twice():
    regular, static = _entry_points()
    dev = torch.device("cuda", torch.cuda.current_device())
    d, c, r, nb, k = 300_007, 40_000, 5, 20, 3_000
    plan = sk.SketchPlan.get(d, c, r, nb, dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    scratch = torch.zeros(int(native.lib().ceb_select_scratch_bytes()), dtype=torch.uint8, device=dev)
    est = torch.empty(d, dtype=torch.float32, device=dev)
    gen = torch.Generator(device=dev)
    for call, seed in enumerate((5, 6)):
        gen.manual_seed(seed)
        g = torch.randn(d, device=dev, generator=gen)
        g[torch.randperm(d, device=dev, generator=gen)[:k]] *= 50.0
        table = plan.new_table()
        sk.accumulate(plan, table, g)

        want_idx = torch.empty(k, dtype=torch.int32, device=dev)
        want_val = torch.zeros(k, dtype=torch.float32, device=dev)
        rc = regular(_p(table), _p(plan.packed), plan.n_tokens, _p(plan.blk), nb, r, c, d, k, _p(est),
                     _p(scratch), scratch.numel(), _p(want_idx), _p(want_val), stream)
        assert rc == 0, native.last_error()

        got_idx = torch.full((k,), -1, dtype=torch.int32, device=dev)
        got_val = torch.zeros(k, dtype=torch.float32, device=dev)
        rc = static(_p(table), _p(plan.packed), plan.n_tokens, _p(plan.blk), nb, r, c, d, k,
                    _p(got_idx), _p(got_val), stream)
        assert rc == 0, native.last_error()
        torch.cuda.synchronize()

        wi, wv = _sorted_by_index(want_idx, want_val)
        gi, gv = _sorted_by_index(got_idx, got_val)
        assert torch.equal(gi, wi), "call %d: different coordinates selected" % call
        torch.testing.assert_close(gv, wv, rtol=0, atol=0)


def test_static_variant_rejects_oversize_and_bad_k():
    _regular, static = _entry_points()
    dev = torch.device("cuda", torch.cuda.current_device())
    dummy = torch.zeros(16, dtype=torch.float32, device=dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    too_big = MAX_D + 1
    rc = static(_p(dummy), _p(dummy), too_big, _p(dummy), 1, 5, 1000, too_big, 10, _p(dummy), _p(dummy), stream)
    assert rc != 0
    msg = native.last_error()
    assert "exceeds the built-in scratch" in msg and str(MAX_D) in msg
    rc = static(_p(dummy), _p(dummy), 100, _p(dummy), 1, 5, 1000, 100, 101, _p(dummy), _p(dummy), stream)
    assert rc != 0 and "k" in native.last_error()
    torch.cuda.synchronize()
