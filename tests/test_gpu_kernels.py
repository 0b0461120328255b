"""Native sm_100a kernels vs plain-PyTorch fp32 references of the same ops."""
import numpy as np
import pytest
import torch

from commefficient_b200.ops import dense as dn
from commefficient_b200.ops import native, ref
from commefficient_b200.ops import sketch as sk
from commefficient_b200.ops.hashing import make_hashes
from commefficient_b200.ops.topk import topk_sparse

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _same_topk(vec, idx, val, k):
    """Selected set must be a valid top-k: right size, unique, values match the vector,
    and the multiset of magnitudes equals torch.topk's."""
    idx = idx.long()
    assert idx.numel() == k and idx.unique().numel() == k
    assert torch.equal(vec[idx], val)
    want = torch.topk(vec.abs(), k).values.sort().values
    got = val.abs().sort().values
    assert torch.equal(want, got)


@pytest.mark.parametrize("n,k", [(1000, 1), (1000, 1000), (4097, 37), (1 << 20, 50000), (6568640, 50000)])
def test_topk_select_random(n, k):
    g = torch.Generator(device=DEV).manual_seed(n + k)
    v = torch.randn(n, device=DEV, generator=g) * torch.rand(n, device=DEV, generator=g)
    idx, val = topk_sparse(v, k)
    _same_topk(v, idx, val, k)


def test_topk_select_ties_zeros_nan():
    v = torch.zeros(5000, device=DEV)
    v[10], v[20], v[30] = 3.0, -3.0, 3.0          # ties at the threshold
    v[40] = float("nan")
    v[50] = float("inf")
    idx, val = topk_sparse(v, 3)
    assert 50 in idx.tolist() and 40 not in idx.tolist()
    assert sorted(val.abs().tolist()) == [3.0, 3.0, float("inf")]
    idx, val = topk_sparse(v, 100)                # more than the non-zeros: padded with zeros
    assert idx.unique().numel() == 100 and int((val != 0).sum()) == 4
    ones = torch.ones(3000, device=DEV)           # all equal
    idx, val = topk_sparse(ones, 17)
    assert idx.unique().numel() == 17 and bool((val == 1).all())


@pytest.mark.parametrize("n,k,nties", [(5000, 40, 100), (1 << 20, 50000, 4000), (300007, 1000, 2500)])
def test_topk_select_ties_are_lowest_index_and_output_is_ordered(n, k, nties):
    """Entries equal to the threshold are taken in ascending index order (deterministic: replicas that
    select on the same vector agree), the output lists the entries above the threshold and then the
    ties, each in ascending index order -- and the CPU oracle follows the same rule."""
    from commefficient_b200.ops.ref import topk_sparse_ref
    g = torch.Generator(device=DEV).manual_seed(n)
    v = torch.rand(n, device=DEV, generator=g) * 0.5
    big = torch.randperm(n, device=DEV, generator=g)[: k - nties // 4]       # strictly above the ties
    v[big] = 2.0 + torch.rand(big.numel(), device=DEV, generator=g)
    tie_pos = torch.randperm(n, device=DEV, generator=g)[:nties]
    tie_pos = tie_pos[~torch.isin(tie_pos, big)]
    v[tie_pos] = 1.0
    v[tie_pos[::2]] = -1.0
    idx, val = topk_sparse(v, k)
    n_above = big.numel()
    want_ties = tie_pos.sort().values[: k - n_above]
    assert torch.equal(idx[:n_above].long(), big.sort().values)
    assert torch.equal(idx[n_above:].long(), want_ties)
    assert torch.equal(val, v[idx.long()])
    for _ in range(3):                                   # and the same answer every time
        idx2, _ = topk_sparse(v, k)
        assert torch.equal(idx, idx2)
    ci, cv = topk_sparse_ref(v.cpu(), k)
    assert torch.equal(ci, idx.cpu().long()) and torch.equal(cv, val.cpu())


@pytest.mark.parametrize("d,c,r,nb", [(1000, 50, 3, 1), (1003, 97, 5, 20), (70001, 5000, 5, 20),
                                       (4096, 16, 1, 4), (50000, 1000, 8, 7), (30000, 999, 16, 3)])
def test_sketch_accumulate_and_query(d, c, r, nb):
    plan = sk.SketchPlan.get(d, c, r, nb, DEV)
    cpu_plan = sk.SketchPlan.get(d, c, r, nb, "cpu")
    g = torch.randn(d, device=DEV)
    w = torch.randn(d, device=DEV)
    table = plan.new_table()
    sk.accumulate(plan, table, g, scale=2.0, weights=w, wd=0.25)
    want = cpu_plan.new_table()
    ref.sketch_accum_ref(want, (g + 0.25 * w).cpu(), cpu_plan.hashes, 2.0)
    torch.testing.assert_close(table.cpu(), want, rtol=1e-4, atol=1e-4)
    est = sk.query(plan, table)
    want_est = ref.sketch_estimate_ref(table.cpu(), cpu_plan.hashes)
    torch.testing.assert_close(est.cpu(), want_est, rtol=0, atol=0)
    sub = torch.randint(0, d, (257,), device=DEV, dtype=torch.int32)
    torch.testing.assert_close(sk.query(plan, table, sub), est[sub.long()], rtol=0, atol=0)


@pytest.mark.parametrize("d,c,r,nb,k", [(20000, 2000, 5, 20, 100), (6001, 601, 4, 1, 6001), (100000, 20000, 5, 20, 5000)])
def test_unsketch_topk(d, c, r, nb, k):
    plan = sk.SketchPlan.get(d, c, r, nb, DEV)
    cpu_plan = sk.SketchPlan.get(d, c, r, nb, "cpu")
    g = torch.randn(d, device=DEV) * 0.01
    heavy = torch.randperm(d, device=DEV)[: max(1, k // 2)]
    g[heavy] += torch.randn(heavy.numel(), device=DEV).sign() * (1 + torch.rand(heavy.numel(), device=DEV))
    table = plan.new_table()
    sk.accumulate(plan, table, g)
    idx, val = sk.unsketch_topk(plan, table, k)
    est = ref.sketch_estimate_ref(table.cpu(), cpu_plan.hashes).to(DEV)
    _same_topk(est, idx, val, k)
    # the heavy coordinates are recovered
    if k < d:
        found = np.isin(heavy.cpu().numpy(), idx.cpu().numpy()).mean()
        assert found > 0.95


def test_l2_scale_server_update_apply_zero():
    d, c, r, nb = 30000, 3000, 5, 20
    plan = sk.SketchPlan.get(d, c, r, nb, DEV)
    cpu_plan = sk.SketchPlan.get(d, c, r, nb, "cpu")
    table = torch.randn(r, c, device=DEV)
    l2 = ref.l2estimate_ref(table.cpu())
    out = sk.l2_scale(plan, table, clip=l2 / 2).clone()
    assert out[1].item() == pytest.approx(l2, rel=1e-5) and out[0].item() == pytest.approx(0.5, rel=1e-5)
    assert sk.l2_scale(plan, table, clip=l2 * 2)[0].item() == 1.0
    y = torch.randn(r, c, device=DEV)
    y0, x0 = y.clone(), table.clone()
    sk.axpy(y, table, alpha=3.0, dev_scale=out, zero_x=True)
    torch.testing.assert_close(y, y0 + 1.5 * x0)
    assert float(table.abs().sum()) == 0.0

    Vv, Ve, g = (torch.randn(r, c, device=DEV) for _ in range(3))
    Vv0, Ve0 = Vv.clone(), Ve.clone()
    sk.server_update(Vv, Ve, g, 0.25, 0.9, "virtual")
    torch.testing.assert_close(Vv, 0.9 * Vv0 + 0.25 * g)
    torch.testing.assert_close(Ve, Ve0 + Vv)

    idx = torch.randperm(d, device=DEV)[:500].to(torch.int32)
    val = torch.randn(500, device=DEV)
    val[:7] = 0                                    # zero-valued selections touch nothing
    w = torch.randn(d, device=DEV)
    w_ref = w.clone()
    Vv_ref, Ve_ref = Vv.cpu().clone(), Ve.cpu().clone()
    sk.apply_zero(plan, idx, val, Vv, Ve, w, 0.1)
    ref.sketch_zero_buckets_ref([Vv_ref, Ve_ref], idx.cpu(), val.cpu(), cpu_plan.hashes)
    w_ref[idx.long()] -= 0.1 * val
    torch.testing.assert_close(Vv.cpu(), Vv_ref, rtol=0, atol=0)
    torch.testing.assert_close(Ve.cpu(), Ve_ref, rtol=0, atol=0)
    torch.testing.assert_close(w, w_ref)
    lr_vec = torch.rand(d, device=DEV)
    w2 = w.clone()
    sk.apply_zero(plan, idx, val, None, None, w2, lr_vec)
    w_ref2 = w.clone()
    w_ref2[idx.long()] -= lr_vec[idx.long()] * val
    torch.testing.assert_close(w2, w_ref2)


def test_dense_server_and_true_topk():
    d, k = 200003, 1000
    Vv, g, w = (torch.randn(d, device=DEV) for _ in range(3))
    Vv0, w0 = Vv.clone(), w.clone()
    dn.dense_server_update(Vv, g, 0.5, 0.9, w, 0.1)
    torch.testing.assert_close(Vv, 0.9 * Vv0 + 0.5 * g)
    torch.testing.assert_close(w, w0 - 0.1 * Vv)
    # server-side DP noise has the requested scale
    Vz, wz = torch.zeros(d, device=DEV), torch.zeros(d, device=DEV)
    dn.dense_server_update(Vz, torch.zeros(d, device=DEV), 1.0, 0.0, wz, 1.0, noise_std=2.0, seed=3)
    assert wz.std().item() == pytest.approx(2.0, rel=0.02) and abs(wz.mean().item()) < 0.05

    Vv, Ve, g, w = (torch.randn(d, device=DEV) for _ in range(4))
    cVv, cVe, cw = Vv.cpu().clone(), Ve.cpu().clone(), w.cpu().clone()
    idx, val = dn.true_topk_step(Vv, Ve, g, 0.25, 0.9, k, w, 0.05)
    ridx, rval = dn.true_topk_step(cVv, cVe, g.cpu(), 0.25, 0.9, k, cw, 0.05)
    assert set(idx.tolist()) == set(ridx.tolist())
    torch.testing.assert_close(Vv.cpu(), cVv)
    torch.testing.assert_close(Ve.cpu(), cVe)
    torch.testing.assert_close(w.cpu(), cw)


@pytest.mark.parametrize("use_vel,use_err,k", [(False, False, 0), (True, False, 0), (True, True, 0),
                                                (False, False, 500), (True, True, 500), (False, True, 500)])
def test_client_step(use_vel, use_err, k):
    d = 100001
    g = torch.randn(d, device=DEV)
    vel = torch.randn(d, device=DEV) if use_vel else None
    err = torch.randn(d, device=DEV) if use_err else None
    acc = torch.randn(d, device=DEV)
    cg, cacc = g.cpu().clone(), acc.cpu().clone()
    cvel = vel.cpu().clone() if use_vel else None
    cerr = err.cpu().clone() if use_err else None
    dn.client_step(g, 5.0, vel, 0.9, err, acc, k=k)
    dn.client_step(cg, 5.0, cvel, 0.9, cerr, cacc, k=k)
    torch.testing.assert_close(acc.cpu(), cacc)
    if use_vel:
        torch.testing.assert_close(vel.cpu(), cvel)
    if use_err:
        torch.testing.assert_close(err.cpu(), cerr)


def test_clip_noise_fedavg_count():
    d = 300007
    x = torch.randn(d, device=DEV)
    norm = x.norm().item()
    y = x.clone()
    dn.clip_by_norm(y, norm / 4, eps=0.0)
    assert y.norm().item() == pytest.approx(norm / 4, rel=1e-4)
    y = x.clone()
    dn.clip_by_norm(y, norm * 4, eps=0.0)
    assert torch.equal(y, x)
    y = x.clone()
    dn.clip_by_norm(y, norm / 2, eps=1e-6)
    assert y.norm().item() == pytest.approx(norm / 2, rel=1e-4)

    z = torch.zeros(d, device=DEV)
    dn.add_gaussian_noise(z, 3.0, seed=11)
    assert z.std().item() == pytest.approx(3.0, rel=0.02) and abs(z.mean().item()) < 0.05
    z2 = torch.zeros(d, device=DEV)
    dn.add_gaussian_noise(z2, 3.0, seed=11)
    assert torch.equal(z, z2)                       # counter-based RNG: reproducible
    z3 = torch.zeros(d, device=DEV)
    dn.add_gaussian_noise(z3, 3.0, seed=11, offset=1 << 20)
    assert not torch.equal(z, z3)

    acc, w0, wl = (torch.randn(d, device=DEV) for _ in range(3))
    want = acc + 7.0 * (w0 - wl)
    dn.fedavg_delta(acc, w0, wl, 7.0)
    torch.testing.assert_close(acc, want)

    w = torch.randn(d, device=DEV)
    prev = w.clone()
    mask = torch.zeros(d, dtype=torch.uint8, device=DEV)
    w[::3] += 1.0
    cnt = dn.count_changed(w, prev, mask)
    assert int(cnt.item()) == len(range(0, d, 3)) and torch.equal(prev, w)
    w[1::3] += 1.0
    cnt = dn.count_changed(w, prev, mask)
    assert int(cnt.item()) == len(range(0, d, 3)) + len(range(1, d, 3))


def test_launch_counter_and_errors():
    native.reset_launch_count()
    topk_sparse(torch.randn(1000, device=DEV), 10)
    assert native.launch_count() == 5          # 3 histogram passes + ordered emit (count, write)
    with pytest.raises(native.NativeError):
        native.call("ceb_topk_select", None, 10, 1, None, 0, None, None, None)


def test_download_tracker_native_matches_torch_path():
    """ceb_tracker_record / ceb_tracker_counts against the torch implementation on the CPU."""
    from commefficient_b200.fed_aggregator import _DownloadTracker
    d, n_clients = 100_003, 40
    dev_tr = _DownloadTracker(d, n_clients, torch.device("cuda"), sparse=True)
    cpu_tr = _DownloadTracker(d, n_clients, torch.device("cpu"), sparse=True)
    g = torch.Generator().manual_seed(3)
    for rnd in range(6):
        clients = np.sort(np.random.RandomState(rnd).choice(n_clients, 7, replace=False))
        a = dev_tr.download_counts(clients)
        b = cpu_tr.download_counts(clients)
        assert a.dtype == torch.float32
        torch.testing.assert_close(a.cpu(), b)
        idx = torch.randperm(d, generator=g)[:5000].to(torch.int32)
        dev_tr.record_sparse(idx.cuda())
        cpu_tr.record_sparse(idx)
        n = cpu_tr.hist.numel()
        assert torch.equal(dev_tr.hist[:n].cpu(), cpu_tr.hist) and int(dev_tr.hist[n:].sum()) == 0
        assert torch.equal(dev_tr.last.cpu(), cpu_tr.last)
    assert int(dev_tr.hist.sum()) == d


def test_sketch_accumulate_range_partitions_sum_to_full():
    from commefficient_b200.ops import sketch as sk
    d, c, r = 1_000_003, 50_000, 5
    plan = sk.SketchPlan.get(d, c, r, 20, "cuda")
    g = torch.randn(d, device="cuda")
    w = torch.randn(d, device="cuda")
    full = plan.new_table()
    sk.accumulate(plan, full, g, scale=0.5, weights=w, wd=0.01)
    parts = plan.new_table()
    cuts = [0, 17, 300_000, 300_001, 777_777, d]
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        sk.accumulate_range(plan, parts, g, lo, hi, scale=0.5, weights=w, wd=0.01)
    torch.testing.assert_close(parts, full, rtol=1e-4, atol=1e-4)
