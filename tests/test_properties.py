"""Property-based tests (hypothesis) of the compression primitives: the algebraic facts the federated rules rely
on, checked on the torch oracles (CPU tier) and -- marked gpu -- on the native kernels against those oracles.
Reference semantics: ``_topk`` (/root/reference/CommEfficient/utils.py:232-252) and the CSVec calls of
fed_worker.py:312-320 / fed_aggregator.py:584-611."""
import numpy as np
import pytest
import torch
from hypothesis import given, settings, strategies as st

from commefficient_b200.ops import ref
from commefficient_b200.ops.hashing import make_hashes

vec = st.integers(min_value=1, max_value=400).flatmap(
    lambda n: st.lists(st.floats(min_value=-100, max_value=100, allow_nan=False, width=32), min_size=n, max_size=n))


@settings(max_examples=60, deadline=None)
@given(vec, st.integers(min_value=1, max_value=500))
def test_topk_oracle_properties(values, k):
    v = torch.tensor(values, dtype=torch.float32)
    k = min(k, v.numel())
    idx, val = ref.topk_sparse_ref(v, k)
    assert idx.numel() == k and idx.unique().numel() == k              # exactly k distinct coordinates
    assert torch.equal(val, v[idx])
    kept = torch.zeros(v.numel(), dtype=torch.bool)
    kept[idx] = True
    if k < v.numel():                                                   # nothing dropped beats anything kept
        assert float(v[~kept].abs().max()) <= float(v[kept].abs().min())
    thresh = v[kept].abs().min()
    ties = torch.nonzero(v.abs() == thresh).flatten()                   # ties are taken in ascending index order
    taken = ties[kept[ties]]
    assert torch.equal(taken, ties[: taken.numel()])
    dense = ref.topk_dense_ref(v, k)
    assert int((dense != 0).sum()) <= k and torch.equal(dense[idx], v[idx])


@settings(max_examples=25, deadline=None)
@given(st.integers(min_value=20, max_value=600), st.integers(min_value=8, max_value=64), st.integers(min_value=1, max_value=5),
       st.integers(min_value=1, max_value=7), st.integers(min_value=0, max_value=2 ** 16))
def test_sketch_is_linear_and_recovers_isolated_heavy_hitters(d, c, r, nb, seed):
    g = torch.Generator().manual_seed(seed)
    h = make_hashes(d, c, r, nb)
    a, b = torch.randn(d, generator=g), torch.randn(d, generator=g)
    ta, tb, tab = (torch.zeros(r, c) for _ in range(3))
    ref.sketch_accum_ref(ta, a, h)
    ref.sketch_accum_ref(tb, b, h)
    ref.sketch_accum_ref(tab, 2.0 * a - 3.0 * b, h)
    torch.testing.assert_close(tab, 2.0 * ta - 3.0 * tb, rtol=1e-4, atol=1e-4)      # S(2a - 3b) = 2 S(a) - 3 S(b)
    # a vector with a single non-zero is estimated exactly, whatever the geometry
    e = torch.zeros(d)
    i = int(torch.randint(0, d, (1,), generator=g))
    e[i] = 7.5
    te = torch.zeros(r, c)
    ref.sketch_accum_ref(te, e, h)
    est = ref.sketch_estimate_ref(te, h)
    assert float(est[i]) == pytest.approx(7.5)
    # heavy-hitter recovery returns it -- or, in a tiny table, a coordinate that collides with it in a majority of
    # the rows and is therefore indistinguishable (same magnitude; the lowest index wins the tie)
    idx, val = ref.unsketch_ref(te, h, 1)
    assert abs(float(val[0])) == pytest.approx(7.5) and int(idx[0]) <= i


@pytest.mark.gpu
@settings(max_examples=30, deadline=None)
@given(st.integers(min_value=1, max_value=70000), st.integers(min_value=1, max_value=5000), st.integers(min_value=0, max_value=2 ** 16),
       st.sampled_from([0, 1, 8, 200]))
def test_native_topk_equals_oracle(n, k, seed, quantise):
    from commefficient_b200.ops.topk import topk_sparse
    g = torch.Generator().manual_seed(seed)
    v = torch.randn(n, generator=g)
    if quantise:
        v = (v * quantise).round() / quantise            # forces ties, including at the threshold and at zero
    k = min(k, n)
    idx, val = topk_sparse(v.cuda(), k)
    ci, cv = ref.topk_sparse_ref(v, k)
    assert torch.equal(idx.cpu().long(), ci) and torch.equal(val.cpu(), cv)


@pytest.mark.gpu
@settings(max_examples=15, deadline=None)
@given(st.integers(min_value=50, max_value=40000), st.integers(min_value=16, max_value=3000), st.integers(min_value=1, max_value=5),
       st.integers(min_value=1, max_value=20), st.integers(min_value=0, max_value=2 ** 16))
def test_native_sketch_equals_oracle(d, c, r, nb, seed):
    from commefficient_b200.ops import sketch as sk
    g = torch.Generator().manual_seed(seed)
    v = torch.randn(d, generator=g)
    plan = sk.SketchPlan.get(d, c, r, nb, "cuda")
    table = plan.new_table()
    sk.accumulate(plan, table, v.cuda())
    want = torch.zeros(r, c)
    ref.sketch_accum_ref(want, v, plan.hashes)
    torch.testing.assert_close(table.cpu(), want, rtol=1e-4, atol=1e-4)
    k = max(1, min(d // 10, 500))
    idx, val = sk.unsketch_topk(plan, table, k)
    est = ref.sketch_estimate_ref(table.cpu(), plan.hashes)
    torch.testing.assert_close(val.cpu(), est[idx.cpu().long()], rtol=1e-5, atol=1e-6)
    thresh = est.abs().topk(k).values[-1]
    assert float(val.abs().min()) >= float(thresh) - 1e-6


class _Sizes:
    def __init__(self, sizes):
        self.data_per_client = list(sizes)
        self.num_clients = len(sizes)

    def __len__(self):
        return int(sum(self.data_per_client))


@settings(max_examples=40, deadline=None)
@given(st.lists(st.integers(min_value=1, max_value=9), min_size=1, max_size=30), st.integers(min_value=1, max_value=8),
       st.sampled_from([-1, 1, 3]), st.integers(min_value=2, max_value=4), st.booleans(), st.integers(min_value=0, max_value=99))
def test_rank_sharded_sampler_partitions_the_unsharded_schedule(sizes, num_workers, local_bs, world, by_owner, seed):
    """For every round the ranks' batches of a sharded FedSampler are a partition of the unsharded sampler's batch
    (or, when some rank would be left without a client, each rank gets the whole batch and the round says so), and
    the published schedule (client ids ascending, sample counts) agrees with the samples drawn."""
    from commefficient_b200.data_utils.fed_sampler import FedSampler, assign_clients
    ds = _Sizes(sizes)
    starts = np.concatenate([[0], np.cumsum(sizes)])
    owner = np.repeat(np.arange(len(sizes)), sizes)
    plain = list(FedSampler(ds, num_workers, local_bs, seed=seed))
    shards = [FedSampler(ds, num_workers, local_bs, seed=seed, rank=r, world_size=world, by_owner=by_owner)
              for r in range(world)]
    per_rank, metas = [], []
    for s in shards:
        per_rank.append(list(s))
        metas.append(list(s.rounds))
    assert all(len(b) == len(plain) for b in per_rank)
    seen = np.zeros(int(sum(sizes)), dtype=np.int64)
    for t, full in enumerate(plain):
        meta = metas[0][t]
        assert all(np.array_equal(m[t].ids, meta.ids) and np.array_equal(m[t].counts, meta.counts) for m in metas)
        assert np.all(np.diff(meta.ids) > 0) and int(meta.counts.sum()) == len(full)
        assert sorted(np.unique(owner[full])) == list(meta.ids)
        parts = assign_clients(meta.ids, world, by_owner)
        if meta.sharded:
            got = np.concatenate([per_rank[r][t] for r in range(world)])
            assert sorted(got) == sorted(full)
            for r in range(world):
                assert sorted(np.unique(owner[per_rank[r][t]])) == list(parts[r])
        else:
            assert any(len(p) == 0 for p in parts)
            assert all(sorted(per_rank[r][t]) == sorted(full) for r in range(world))
        seen[full] += 1
    assert np.all(seen == 1) and starts[-1] == len(seen)          # an epoch visits every sample exactly once


@settings(max_examples=60, deadline=None)
@given(st.integers(min_value=1, max_value=200), st.integers(min_value=1, max_value=40), st.integers(min_value=1, max_value=9),
       st.integers(min_value=1, max_value=5))
def test_sharded_eval_sampler_partitions_every_batch(num_items, batch_size, shard_size, world):
    """Per validation batch the ranks' index lists are, in rank order, exactly the sequential batch (or, when some rank
    would get no shard, every rank gets the whole batch and the entry says so); one rank = the plain sequential loader."""
    from commefficient_b200.data_utils.fed_sampler import ShardedEvalSampler
    plain = list(ShardedEvalSampler(num_items, batch_size, shard_size))
    assert [i for b in plain for i in b] == list(range(num_items))
    assert all(len(b) == batch_size for b in plain[:-1]) and len(plain) == -(-num_items // batch_size)
    samplers = [ShardedEvalSampler(num_items, batch_size, shard_size, r, world) for r in range(world)]
    per_rank = [list(s) for s in samplers]
    for t, full in enumerate(plain):
        if world == 1:
            assert per_rank[0][t] == full
            continue
        meta = samplers[0].rounds[t]
        assert meta.total == len(full) and all(s.rounds[t] == meta for s in samplers)
        if meta.sharded:
            assert [i for r in range(world) for i in per_rank[r][t]] == full
            assert all(len(per_rank[r][t]) > 0 for r in range(world))
        else:
            assert all(per_rank[r][t] == full for r in range(world))
