"""Multi-rank tests: gloo on CPU (host-side protocol; BASELINE.json config 1) and, on a box
with >= 2 GPUs, the peer-memory kernels against NCCL."""
import os
import tempfile

import numpy as np
import pytest
import torch

from commefficient_b200.parallel import runtime

SMALL = {"prep": 8, "layer1": 16, "layer2": 16, "layer3": 32}


def _fed_run(device, mode, backend, rounds=3, **kw):
    from commefficient_b200 import models
    from commefficient_b200.cv_train import compute_loss_train, compute_loss_val
    from commefficient_b200.data_utils import synthetic_image_batch
    from commefficient_b200.fed_aggregator import FedModel, FedOptimizer
    from commefficient_b200.utils import make_args
    torch.manual_seed(0)
    base = dict(mode=mode, num_workers=6, num_clients=36, local_batch_size=-1, k=300, num_cols=4000, num_rows=5,
                num_blocks=4, device=device, dataset_name="CIFAR10", precision="fp32", weight_decay=5e-4,
                local_momentum=0.0, virtual_momentum=0.9, flat_layout="reference", backend=backend,
                error_type="virtual" if mode in ("sketch", "true_topk") else "none")
    base.update(kw)
    bn = base.pop("bn", False)
    args = make_args(**base)
    model = models.ResNet9(num_classes=10, channels=SMALL, do_batchnorm=bn)
    fed = FedModel(model, compute_loss_train, args, compute_loss_val)
    opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args)
    opt.param_groups[0]["lr"] = 0.05
    losses = []
    for step in range(rounds):
        fed.train(True)
        out = fed(synthetic_image_batch(6, 4, seed=step, first_client=6 * step))
        opt.step()
        losses.append([float(x) for x in out[0]])
    fed.train(False)
    val = fed(synthetic_image_batch(4, 8, seed=99))
    w = fed.fp.to_reference_order(fed.fp.flat).detach().cpu()
    return w, losses, [float(x) for x in val[0]], fed


def _rank_main(rank, world, device, mode, backend, out_dir, kw):
    rt = runtime.init_runtime(rank=rank, world_size=world, device=device, backend=backend)
    w, losses, val, fed = _fed_run(device, mode, backend, **kw)
    torch.save({"w": w, "losses": losses, "val": val, "backend": rt.backend}, os.path.join(out_dir, "r%d.pt" % rank))
    rt.shutdown()


def _spawn(world, device, mode, backend, **kw):
    out_dir = tempfile.mkdtemp()
    runtime.spawn(_rank_main, world, args=(device, mode, backend, out_dir, kw), device=device)
    return [torch.load(os.path.join(out_dir, "r%d.pt" % r)) for r in range(world)]


@pytest.mark.parametrize("mode,kw", [("sketch", {}), ("uncompressed", {}), ("true_topk", {}),
                                     ("local_topk", {"error_type": "local", "local_momentum": 0.9}),
                                     ("fedavg", {})])
def test_gloo_two_ranks_match_single_rank(mode, kw):
    ranks = _spawn(2, "cpu", mode, "gloo", **kw)
    assert ranks[0]["backend"] == "gloo"
    torch.testing.assert_close(ranks[0]["w"], ranks[1]["w"], rtol=0, atol=0)      # replicas stay identical
    assert ranks[0]["losses"] == ranks[1]["losses"]
    w1, l1, v1, fed = _fed_run("cpu", mode, "auto", **kw)
    fed.rt.shutdown()
    np.testing.assert_allclose(ranks[0]["losses"], l1, rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(ranks[0]["val"], v1, rtol=1e-4, atol=1e-5)
    assert (ranks[0]["w"] - w1).norm() / w1.norm() < 1e-3


def test_gloo_two_ranks_batchnorm_model_matches_single_rank():
    """A BatchNorm model on two ranks: every rank batches its own clients with segment-wise BN statistics;
    the result equals the single-rank run (which batches all six clients) and the replicas stay identical."""
    ranks = _spawn(2, "cpu", "uncompressed", "gloo", bn=True)
    torch.testing.assert_close(ranks[0]["w"], ranks[1]["w"], rtol=0, atol=0)
    w1, l1, _v1, fed = _fed_run("cpu", "uncompressed", "auto", bn=True)
    assert fed.worker.has_bn and fed.worker.can_batch_clients([4] * 6)
    fed.rt.shutdown()
    np.testing.assert_allclose(ranks[0]["losses"], l1, rtol=1e-4, atol=1e-5)
    assert (ranks[0]["w"] - w1).norm() / w1.norm() < 1e-3


def test_gloo_three_ranks_uneven_clients():
    # 6 clients over 3 ranks with per-client state: owner = client id mod 3; nothing is dropped (Q1)
    ranks = _spawn(3, "cpu", "local_topk", "gloo", error_type="local", local_momentum=0.9)
    for r in ranks[1:]:
        torch.testing.assert_close(ranks[0]["w"], r["w"], rtol=0, atol=0)
    assert len(ranks[0]["losses"][0]) == 6


def _need_gpus(n):
    if not torch.cuda.is_available() or torch.cuda.device_count() < n:
        pytest.skip("needs %d GPUs" % n)


def _peer_kernel_main(rank, world, out_dir):
    rt = runtime.init_runtime(rank=rank, world_size=world, device="cuda", backend="b200")
    assert rt.backend == "b200" and rt.peer is not None, "peer backend did not come up"
    peer, dev = rt.peer, rt.device
    gen = torch.Generator(device=dev).manual_seed(100 + rank)
    ok = {}
    # all-reduce of an arbitrary tensor (staged) against NCCL
    for n in (5, 4096, 1_000_003):
        x = torch.randn(n, device=dev, generator=gen)
        want = x.clone()
        torch.distributed.all_reduce(want)
        got = peer.all_reduce_sum_(x.clone())
        ok["allreduce_%d" % n] = float((got - want).abs().max())
    # fused sketch: reduce-scatter + momentum/error + all-gather
    n = 5 * 4000
    tr = peer.symm_empty(peer.padded(n))
    ve = peer.symm_empty(peer.padded(n))
    tr.tensor[:n].copy_(torch.randn(n, device=dev, generator=gen))
    total = tr.tensor[:n].clone()
    torch.distributed.all_reduce(total)
    g2 = torch.Generator(device=dev).manual_seed(7)          # same state on every rank
    Vv = torch.randn(n + 4, device=dev, generator=g2)[:n]
    ve.tensor[:n].copy_(torch.randn(n, device=dev, generator=g2))
    Vv_ref = 0.9 * Vv + total * 0.25
    Ve_ref = ve.tensor[:n] + Vv_ref
    lo, hi = peer.shard_range(peer.padded(n))
    peer.fused_sketch(tr, ve, Vv[lo:], ve.tensor[lo:], 0.25, 0.9, "virtual")
    torch.cuda.synchronize()
    ok["sketch_Ve"] = float((ve.tensor[:n] - Ve_ref).abs().max())
    ok["sketch_Vv_shard"] = float((Vv[lo:min(hi, n)] - Vv_ref[lo:min(hi, n)]).abs().max())
    # fused dense: reduce-scatter + momentum + apply + all-gather of weights
    n = 100_003
    tr2 = peer.symm_empty(peer.padded(n))
    wb = peer.symm_empty(peer.padded(n))
    tr2.tensor[:n].copy_(torch.randn(n, device=dev, generator=gen))
    total = tr2.tensor[:n].clone()
    torch.distributed.all_reduce(total)
    Vv = torch.randn(peer.padded(n) + 4, device=dev, generator=g2)[:peer.padded(n)]
    wb.tensor[:n].copy_(torch.randn(n, device=dev, generator=g2))
    Vv_ref = 0.9 * Vv[:n] + total * 0.5
    w_ref = wb.tensor[:n] - 0.1 * Vv_ref
    lo, hi = peer.shard_range(peer.padded(n))
    peer.fused_dense(tr2, wb, Vv[lo:], 0.5, 0.9, 0.1)
    torch.cuda.synchronize()
    ok["dense_w"] = float((wb.tensor[:n] - w_ref).abs().max())
    # the NVLS variant (multimem.ld_reduce / multimem.st through the switch) against the peer-load variant on the
    # same buffers; where the box has no multicast support the group reports nvls = False and only the latter runs
    ok["info_nvls"] = 1.0 if peer.nvls else 0.0
    if peer.nvls:
        assert tr2.mc_ptr and wb.mc_ptr
        n = 100_003
        src = torch.randn(n, device=dev, generator=gen)
        w0 = torch.randn(n, device=dev, generator=g2)
        v0 = torch.randn(peer.padded(n) + 4, device=dev, generator=g2)[:peer.padded(n)]
        outs = []
        default = peer.use_nvls
        for use_nvls in (True, False):
            peer.use_nvls = use_nvls
            tr2.tensor[:n].copy_(src)
            wb.tensor[:n].copy_(w0)
            Vv = v0.clone()
            torch.cuda.synchronize()
            torch.distributed.barrier()
            peer.fused_dense(tr2, wb, Vv[lo:], 0.5, 0.9, 0.1)
            torch.cuda.synchronize()
            torch.distributed.barrier()
            outs.append(wb.tensor[:n].clone())
        peer.use_nvls = default
        ok["nvls_vs_p2p_dense"] = float((outs[0] - outs[1]).abs().max())
    # distributed exact top-k over per-rank slices (with ties at the threshold) against the single-GPU select
    from commefficient_b200.ops import native
    from commefficient_b200.ops.topk import topk_sparse
    from commefficient_b200.parallel.symm import DistSelect
    for n_total, k, nties in ((100_003, 1000, 0), (1_000_000, 50_000, 3000), (5000, 4999, 0)):
        g3 = torch.Generator(device=dev).manual_seed(n_total)              # the same full vector on every rank
        full = torch.randn(n_total, device=dev, generator=g3)
        if nties:
            thr = full.abs().topk(k).values[-1]
            pos = torch.randperm(n_total, device=dev, generator=g3)[:nties]
            full[pos] = thr * torch.where(torch.rand(nties, device=dev, generator=g3) < 0.5, -1.0, 1.0)
        lo, hi = DistSelect.coord_range(n_total, rank, world)
        ds = DistSelect(peer, k)
        for rep in range(2):                                               # second call: fresh epoch, reused buffers
            idx, val = ds.run(full[lo:hi].contiguous(), lo, native.select_scratch(dev), pass1_done=False)
            torch.cuda.synchronize()
            torch.distributed.barrier()
            want_idx, want_val = topk_sparse(full, k)
            same = torch.equal(idx.long().sort().values, want_idx.long().sort().values)
            ok["dsel_%d_%d_set" % (n_total, rep)] = 0.0 if same else 1.0
            ok["dsel_%d_%d_val" % (n_total, rep)] = float((full[idx.long()] - val).abs().max())
            gathered = [torch.empty_like(idx) for _ in range(world)]
            torch.distributed.all_gather(gathered, idx.contiguous())
            ok["dsel_%d_%d_replicas" % (n_total, rep)] = 0.0 if all(torch.equal(gathered[0], t) for t in gathered) else 1.0
    peer.check()
    torch.save(ok, os.path.join(out_dir, "r%d.pt" % rank))
    rt.shutdown()


@pytest.mark.gpu
def test_peer_kernels_two_gpus():
    _need_gpus(2)
    out_dir = tempfile.mkdtemp()
    runtime.spawn(_peer_kernel_main, 2, args=(out_dir,), device="cuda")
    for r in range(2):
        res = torch.load(os.path.join(out_dir, "r%d.pt" % r))
        for name, err in res.items():
            if name.startswith("info_"):
                print("rank %d %s = %s" % (r, name, err))
                continue
            assert err < 1e-4, (r, name, err)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["sketch", "true_topk"])
def test_sharded_server_tail_equals_replicated_tail(mode):
    """The sharded tail (each rank estimates / owns half the coordinates, distributed select) picks the
    coordinates the replicated tail picks: replicas identical bit for bit within a run, and the two runs agree to
    the run-to-run noise of the (library, small-model) convolutions -- one differently chosen coordinate would
    move the weights by three orders of magnitude more."""
    _need_gpus(2)
    a = _spawn(2, "cuda", mode, "b200", rounds=4)
    b = _spawn(2, "cuda", mode, "b200", rounds=4, replicated_tail=True)
    torch.testing.assert_close(a[0]["w"], a[1]["w"], rtol=0, atol=0)
    torch.testing.assert_close(b[0]["w"], b[1]["w"], rtol=0, atol=0)
    np.testing.assert_allclose(a[0]["losses"], b[0]["losses"], rtol=1e-5, atol=1e-5)
    assert float((a[0]["w"] - b[0]["w"]).norm() / b[0]["w"].norm()) < 1e-5


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["sketch", "uncompressed", "true_topk"])
def test_peer_backend_matches_nccl_backend(mode):
    _need_gpus(2)
    a = _spawn(2, "cuda", mode, "b200")
    b = _spawn(2, "cuda", mode, "nccl")
    assert a[0]["backend"] == "b200" and b[0]["backend"] == "nccl"
    torch.testing.assert_close(a[0]["w"], a[1]["w"], rtol=0, atol=0)            # replicas identical
    np.testing.assert_allclose(a[0]["losses"], b[0]["losses"], rtol=2e-3, atol=2e-3)
    assert (a[0]["w"] - b[0]["w"]).norm() / b[0]["w"].norm() < (5e-2 if mode != "uncompressed" else 1e-3)


# --------------------------------------------------------------------------- resume sidecar on two ranks
def _resume_main(rank, world, out_dir, phase):
    from torch.optim.lr_scheduler import LambdaLR

    from commefficient_b200 import models
    from commefficient_b200.cv_train import compute_loss_train, compute_loss_val
    from commefficient_b200.data_utils import synthetic_image_batch
    from commefficient_b200.fed_aggregator import FedModel, FedOptimizer
    from commefficient_b200.utils import make_args
    rt = runtime.init_runtime(rank=rank, world_size=world, device="cpu", backend="gloo")
    torch.manual_seed(0)
    args = make_args(mode="local_topk", num_workers=4, num_clients=8, local_batch_size=-1, k=300, device="cpu",
                     dataset_name="CIFAR10", precision="fp32", weight_decay=5e-4, local_momentum=0.9,
                     virtual_momentum=0.9, flat_layout="reference", backend="gloo", error_type="local")
    model = models.ResNet9(num_classes=10, channels=SMALL)
    fed = FedModel(model, compute_loss_train, args, compute_loss_val)
    opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args)
    path = os.path.join(out_dir, "state.fedstate.pt")
    if phase == "resume":
        fed.load_fedstate(path, opt)
    sched = LambdaLR(opt, lr_lambda=lambda step: 0.01 * (step + 1))
    if phase == "resume":
        fed.restore_scheduler(sched)

    def rnd(step):
        sched.step()
        fed.train(True)
        out = fed(synthetic_image_batch(4, 3, seed=step, first_client=(4 * step) % 8))
        opt.step()
        return float(np.mean(out[0]))

    losses = []
    if phase == "straight":
        for s in range(2):
            rnd(s)
        fed.save_fedstate(path, opt, sched)           # every rank writes <path>.rank<r>
        losses = [rnd(s) for s in (2, 3)]
    else:
        losses = [rnd(s) for s in (2, 3)]
    torch.save({"losses": losses, "w": fed.fp.flat.clone(), "lr": opt.param_groups[0]["lr"]},
               os.path.join(out_dir, "%s_r%d.pt" % (phase, rank)))
    rt.shutdown()


def test_resume_sidecar_two_ranks_roundtrip():
    """Every rank saves and loads its own sidecar (client-state rows owned by client id mod world size, the
    scheduler step, the flat layout): the resumed two-rank run continues exactly like the uninterrupted one."""
    out_dir = tempfile.mkdtemp()
    runtime.spawn(_resume_main, 2, args=(out_dir, "straight"), device="cpu")
    assert all(os.path.exists(os.path.join(out_dir, "state.fedstate.pt.rank%d" % r)) for r in range(2))
    runtime.spawn(_resume_main, 2, args=(out_dir, "resume"), device="cpu")
    for r in range(2):
        a = torch.load(os.path.join(out_dir, "straight_r%d.pt" % r))
        b = torch.load(os.path.join(out_dir, "resume_r%d.pt" % r))
        assert a["losses"] == pytest.approx(b["losses"], rel=1e-6)
        torch.testing.assert_close(a["w"], b["w"])
        assert a["lr"] == pytest.approx(b["lr"]) and a["lr"] == pytest.approx(0.01 * 5)


def test_resume_sidecar_rejects_other_layout(tmp_path):
    from tests.test_features_cpu import _make, _round
    fed, opt, _ = _make("sketch")
    _round(fed, opt, 0)
    path = str(tmp_path / "s.pt")
    fed.save_fedstate(path, opt)
    st = torch.load(path, weights_only=False)
    st["flat_layout"] = "nhwc"
    torch.save(st, path)
    with pytest.raises(ValueError, match="flat_layout"):
        fed.load_fedstate(path, opt)
    fed.rt.shutdown()


# --------------------------------------------------------------------------- DP noise is independent per client
def _dp_main(rank, world, out_dir):
    from commefficient_b200 import fed_worker
    from commefficient_b200.ops import dense as dn
    rt = runtime.init_runtime(rank=rank, world_size=world, device="cpu", backend="gloo")
    # the j-th client of every rank: same position in the rank's loop, different client id
    x = torch.zeros(4096)
    dn.add_gaussian_noise(x, 1.0, seed=fed_worker.noise_seed(5, client_id=10 + rank, round_no=3))
    torch.save(x, os.path.join(out_dir, "n%d.pt" % rank))
    rt.shutdown()


def test_dp_noise_differs_across_ranks_and_rounds():
    from commefficient_b200 import fed_worker
    out_dir = tempfile.mkdtemp()
    runtime.spawn(_dp_main, 2, args=(out_dir,), device="cpu")
    a, b = (torch.load(os.path.join(out_dir, "n%d.pt" % r)) for r in range(2))
    assert float((a - b).abs().max()) > 0.1 and abs(float(torch.corrcoef(torch.stack([a, b]))[0, 1])) < 0.1
    assert float((a - b).std()) == pytest.approx(2 ** 0.5, rel=0.1)      # the difference does not cancel
    seeds = {fed_worker.noise_seed(5, c, r) for c in range(50) for r in range(50)}
    assert len(seeds) == 2500


# ------------------------------------------------------------------ rank-sharded data loading
class _CountingClients(torch.utils.data.Dataset):
    """12 clients x 5 samples of 3x32x32 noise; remembers which samples were actually loaded."""

    def __init__(self):
        g = torch.Generator().manual_seed(7)
        self.data_per_client = [5] * 12
        self.num_clients = 12
        self.x = torch.randn(60, 3, 32, 32, generator=g)
        self.y = torch.randint(0, 10, (60,), generator=g)
        self.loaded = []

    def __len__(self):
        return 60

    def __getitem__(self, i):
        i = int(i)
        self.loaded.append(i)
        return i // 5, self.x[i], self.y[i]


def _sharded_loader_run(device, backend, attach, workers=4):
    from torch.utils.data import DataLoader
    from commefficient_b200 import models
    from commefficient_b200.cv_train import compute_loss_train, compute_loss_val
    from commefficient_b200.data_utils import FedSampler
    from commefficient_b200.fed_aggregator import FedModel, FedOptimizer
    from commefficient_b200.utils import make_args
    torch.manual_seed(0)
    args = make_args(mode="sketch", num_workers=workers, num_clients=12, local_batch_size=3, k=300, num_cols=4000, num_rows=5,
                     num_blocks=4, device=device, dataset_name="CIFAR10", precision="fp32", weight_decay=5e-4,
                     local_momentum=0.0, virtual_momentum=0.9, flat_layout="reference", backend=backend,
                     error_type="virtual")
    model = models.ResNet9(num_classes=10, channels=SMALL, do_batchnorm=False)
    fed = FedModel(model, compute_loss_train, args, compute_loss_val)
    opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args)
    opt.param_groups[0]["lr"] = 0.05
    ds = _CountingClients()
    sampler = FedSampler(ds, args.num_workers, args.local_batch_size, seed=3)
    loader = DataLoader(ds, batch_sampler=sampler, num_workers=0)
    if attach:
        fed.attach_sampler(sampler)
    losses, rounds = [], 0
    fed.train(True)
    for batch in loader:
        out = fed(batch)
        opt.step()
        losses.append(sorted(float(x) for x in out[0]))
        rounds += 1
    w = fed.fp.to_reference_order(fed.fp.flat).detach().cpu()
    return w, losses, sorted(ds.loaded), fed


def _sharded_rank_main(rank, world, out_dir, workers=4):
    rt = runtime.init_runtime(rank=rank, world_size=world, device="cpu", backend="gloo")
    w, losses, loaded, _fed = _sharded_loader_run("cpu", "gloo", attach=True, workers=workers)
    torch.save({"w": w, "losses": losses, "loaded": loaded}, os.path.join(out_dir, "r%d.pt" % rank))
    rt.shutdown()


def test_rank_sharded_loader_loads_each_sample_once_and_matches_single_rank():
    """With ``attach_sampler`` each rank's DataLoader materialises only its own clients (the two ranks' loaded
    sets partition the epoch) and the run equals the single-rank one."""
    out_dir = tempfile.mkdtemp()
    runtime.spawn(_sharded_rank_main, 2, args=(out_dir,), device="cpu")
    r0, r1 = (torch.load(os.path.join(out_dir, "r%d.pt" % r)) for r in range(2))
    torch.testing.assert_close(r0["w"], r1["w"], rtol=0, atol=0)
    assert r0["losses"] == r1["losses"]
    assert sorted(r0["loaded"] + r1["loaded"]) == list(range(60)), "every sample loaded by exactly one rank"
    assert 20 <= len(r0["loaded"]) <= 40
    w1, l1, loaded1, fed = _sharded_loader_run("cpu", "auto", attach=False)
    fed.rt.shutdown()
    assert loaded1 == list(range(60))
    np.testing.assert_allclose(r0["losses"], l1, rtol=1e-4, atol=1e-5)
    assert (r0["w"] - w1).norm() / w1.norm() < 1e-3


def _replica_check_main(rank, world, out_dir):
    rt = runtime.init_runtime(rank=rank, world_size=world, device="cpu", backend="gloo")
    torch.manual_seed(0)
    t = torch.randn(100_003)
    same = rt.replicas_agree(t)
    if rank == 1:
        t[77_777] = torch.nextafter(t[77_777], torch.tensor(10.0))       # one ulp on one rank
    one_ulp = rt.replicas_agree(t)
    u = torch.randn(1000, generator=torch.Generator().manual_seed(5))
    if rank == 1:
        u = u.flip(0).contiguous().flip(0).contiguous()                   # same values
        u[[3, 4]] = u[[4, 3]]                                             # ... two of them swapped
    swapped = rt.replicas_agree(u)
    torch.save({"same": same, "one_ulp": one_ulp, "swapped": swapped}, os.path.join(out_dir, "r%d.pt" % rank))
    rt.shutdown()


def test_replica_checksum_sees_one_ulp_and_permutations():
    out_dir = tempfile.mkdtemp()
    runtime.spawn(_replica_check_main, 2, args=(out_dir,), device="cpu")
    for r in range(2):
        res = torch.load(os.path.join(out_dir, "r%d.pt" % r))
        assert res["same"] is True and res["one_ulp"] is False and res["swapped"] is False


def test_gloo_two_ranks_with_replica_check_every_round():
    """--replica_check_every 1: the check runs before every round and passes (the replicas of a correct run are
    bit-identical)."""
    ranks = _spawn(2, "cpu", "sketch", "gloo", replica_check_every=1)
    torch.testing.assert_close(ranks[0]["w"], ranks[1]["w"], rtol=0, atol=0)


@pytest.mark.parametrize("world,mode,kw", [(4, "sketch", {}), (8, "sketch", {}),
                                           (8, "local_topk", {"error_type": "local", "local_momentum": 0.9})])
def test_gloo_more_ranks_than_clients_per_rank(world, mode, kw):
    """Six clients a round over four ranks (2, 2, 1, 1) and over eight ranks (two of them idle; with per-client state
    the owner rule leaves ranks 6 and 7 without a client): replicas stay identical and equal the one-rank run."""
    ranks = _spawn(world, "cpu", mode, "gloo", **kw)
    for r in ranks[1:]:
        torch.testing.assert_close(ranks[0]["w"], r["w"], rtol=0, atol=0)
        assert r["losses"] == ranks[0]["losses"]
    w1, l1, v1, fed = _fed_run("cpu", mode, "auto", **kw)
    fed.rt.shutdown()
    np.testing.assert_allclose(ranks[0]["losses"], l1, rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(ranks[0]["val"], v1, rtol=1e-4, atol=1e-5)
    assert (ranks[0]["w"] - w1).norm() / w1.norm() < 1e-5


def test_rank_sharded_loader_with_one_client_per_round():
    """One client per round on two ranks: sharding would leave a rank without a batch, so the sampler hands every rank
    the whole round (and says so in the schedule); the run still equals the single-rank one."""
    out_dir = tempfile.mkdtemp()
    runtime.spawn(_sharded_rank_main, 2, args=(out_dir, 1), device="cpu")
    r0, r1 = (torch.load(os.path.join(out_dir, "r%d.pt" % r)) for r in range(2))
    torch.testing.assert_close(r0["w"], r1["w"], rtol=0, atol=0)
    assert r0["loaded"] == list(range(60)) and r1["loaded"] == list(range(60))
    w1, l1, _loaded, fed = _sharded_loader_run("cpu", "auto", attach=False, workers=1)
    fed.rt.shutdown()
    np.testing.assert_allclose(r0["losses"], l1, rtol=1e-4, atol=1e-5)
    assert (r0["w"] - w1).norm() / w1.norm() < 1e-3
