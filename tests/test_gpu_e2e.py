"""End-to-end rounds on the GPU (native kernels) against the same rounds on the CPU
(torch oracles)."""
import numpy as np
import pytest
import torch

from commefficient_b200 import models
from commefficient_b200.cv_train import compute_loss_train, compute_loss_val
from commefficient_b200.data_utils import synthetic_image_batch
from commefficient_b200.fed_aggregator import FedModel, FedOptimizer
from commefficient_b200.ops import native
from commefficient_b200.utils import make_args

pytestmark = pytest.mark.gpu

SMALL = {"prep": 8, "layer1": 16, "layer2": 16, "layer3": 32}


def _run(device, mode, rounds=3, **kw):
    torch.manual_seed(0)
    base = dict(mode=mode, num_workers=4, num_clients=32, local_batch_size=-1, k=300, num_cols=4000, num_rows=5,
                num_blocks=4, device=device, dataset_name="CIFAR10", precision="fp32", weight_decay=5e-4,
                local_momentum=0.0, virtual_momentum=0.9, flat_layout="reference",
                error_type="virtual" if mode in ("sketch", "true_topk") else "none")
    base.update(kw)
    args = make_args(**base)
    model = models.ResNet9(num_classes=10, channels=SMALL)
    fed = FedModel(model, compute_loss_train, args, compute_loss_val)
    opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args)
    opt.param_groups[0]["lr"] = 0.05
    losses = []
    for step in range(rounds):
        fed.train(True)
        out = fed(synthetic_image_batch(4, 6, seed=step, first_client=4 * step))
        opt.step()
        losses.append(float(np.mean(out[0])))
    w = fed.fp.to_reference_order(fed.fp.flat).detach().cpu()
    fed.rt.shutdown()
    return w, losses, out


@pytest.mark.parametrize("mode,kw", [
    ("uncompressed", {}),
    ("uncompressed", {"local_momentum": 0.9}),
    ("sketch", {}),
    ("sketch", {"max_grad_norm": 1.0}),
    ("true_topk", {}),
    ("local_topk", {"error_type": "local", "local_momentum": 0.9}),
    ("fedavg", {"num_fedavg_epochs": 2, "fedavg_batch_size": 3}),
])
def test_round_matches_cpu(mode, kw):
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        native.reset_launch_count()
        wg, lg, out = _run("cuda", mode, **kw)
        assert native.launch_count() > 0, "native kernels did not run"
        wc, lc, _ = _run("cpu", mode, **kw)
    finally:
        torch.backends.cudnn.allow_tf32 = old
    np.testing.assert_allclose(lg, lc, rtol=2e-3, atol=2e-3)
    # sparse modes may pick different near-tie coordinates; compare in aggregate
    rel = (wg - wc).norm() / wc.norm()
    assert rel < (5e-2 if mode in ("sketch", "true_topk", "local_topk") else 1e-3), rel
    assert out[-1].sum() > 0 and out[-2].shape == out[-1].shape


def test_client_batching_equals_per_client():
    w1, l1, _ = _run("cuda", "sketch")
    w2, l2, _ = _run("cuda", "sketch", no_client_batching=True)
    np.testing.assert_allclose(l1, l2, rtol=1e-3, atol=1e-3)
    assert (w1 - w2).norm() / w2.norm() < 5e-2


def test_nhwc_flat_layout_is_a_relabelling():
    """Storing conv weights (out,kh,kw,in) changes no maths: dense modes give the same weights."""
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        w1, l1, _ = _run("cuda", "uncompressed", flat_layout="reference")
        w2, l2, _ = _run("cuda", "uncompressed", flat_layout="nhwc")
    finally:
        torch.backends.cudnn.allow_tf32 = old
    np.testing.assert_allclose(l1, l2, rtol=1e-4, atol=1e-4)
    assert (w1 - w2).norm() / w1.norm() < 1e-4


def test_smoke_entry():
    import __graft_entry__ as g
    g.smoke()


def test_deferred_results_equal_synchronous_results():
    """--defer_results: same weights, losses, accuracies and byte counts as the blocking path; the
    values may be read after opt.step() (that is the point)."""
    w_a, loss_a, out_a = _run("cuda", "sketch", rounds=4)
    w_b, loss_b, out_b = _run("cuda", "sketch", rounds=4, defer_results=True)
    torch.testing.assert_close(w_a, w_b, rtol=1e-5, atol=1e-6)
    assert loss_a == pytest.approx(loss_b, rel=1e-5)
    np.testing.assert_allclose(np.asarray(out_a[1]), np.asarray(out_b[1]))
    total = torch.zeros(32)
    total += out_b[2]
    torch.testing.assert_close(total, out_a[2])
    assert float(out_b[3].sum()) == float(out_a[3].sum())


def test_sketch_overlap_equals_sequential_sketch():
    """Sketching finished gradient ranges on a second stream during the backward pass (native ResNet-9,
    nhwc layout) gives the same weights as sketching after it."""
    full = {"prep": 64, "layer1": 64, "layer2": 64, "layer3": 64}

    def run(**kw):
        torch.manual_seed(0)
        args = make_args(mode="sketch", num_workers=4, num_clients=32, local_batch_size=-1, k=3000, num_cols=40000,
                         num_rows=5, num_blocks=4, device="cuda", dataset_name="CIFAR10", weight_decay=5e-4,
                         local_momentum=0.0, virtual_momentum=0.9, error_type="virtual", **kw)
        model = models.ResNet9(num_classes=10, channels=full)
        fed = FedModel(model, compute_loss_train, args, compute_loss_val)
        opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args)
        opt.param_groups[0]["lr"] = 0.05
        tables = []
        for step in range(5):           # rounds 3+ replay the CUDA graph
            fed.train(True)
            fed(synthetic_image_batch(4, 6, seed=step, first_client=4 * step))
            tables.append(fed.worker.transmit.clone())
            opt.step()
        w = fed.fp.flat.clone()
        fed.rt.shutdown()
        return w, tables

    w_a, t_a = run(no_sketch_overlap=True)
    w_b, t_b = run()
    for a, b in zip(t_a, t_b):
        torch.testing.assert_close(a, b, rtol=1e-3, atol=1e-3)
    assert float((w_a - w_b).abs().max()) < 5e-3


def test_batchnorm_model_batched_graph_round_on_gpu():
    """ResNet-9 with BatchNorm, uncompressed mode: the batched round (segment-wise BN, CUDA graph from the
    third round on; round 4 has unequal client sizes) tracks the one-backward-per-client path."""
    def run(**kw):
        torch.manual_seed(0)
        args = make_args(mode="uncompressed", num_workers=4, num_clients=32, local_batch_size=-1, device="cuda",
                         dataset_name="CIFAR10", precision="fp32", weight_decay=5e-4, local_momentum=0.0,
                         virtual_momentum=0.9, error_type="none", flat_layout="reference", **kw)
        model = models.ResNet9(num_classes=10, channels=SMALL, do_batchnorm=True)
        fed = FedModel(model, compute_loss_train, args, compute_loss_val)
        opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args)
        opt.param_groups[0]["lr"] = 0.05
        losses = []
        for step in range(5):
            fed.train(True)
            batch = synthetic_image_batch(4, 6, seed=step, first_client=4 * step)
            if step == 4:
                keep = torch.tensor([i for i in range(24) if i not in (7, 20, 21)])
                batch = tuple(t[keep] for t in batch)
            out = fed(batch)
            opt.step()
            losses.append(float(np.mean(out[0])))
        w = fed.fp.flat.detach().cpu().clone()
        batched = fed.worker.can_batch_clients([6, 6, 6, 6])
        fed.rt.shutdown()
        return w, losses, batched

    w_b, l_b, batched = run()
    w_c, l_c, per_client = run(no_client_batching=True)
    assert batched and not per_client
    assert l_b == pytest.approx(l_c, rel=2e-3, abs=2e-3)
    # (TF32 convolutions pick different algorithms for the two batch shapes and 6-sample BatchNorm amplifies
    # the rounding; the exact equivalence is asserted in fp32 on the CPU, tests/test_features_cpu.py)
    assert float((w_b - w_c).norm() / w_c.norm()) < 1e-2


@pytest.mark.parametrize("mode,kw", [("sketch", {}), ("local_topk", {"error_type": "local", "local_momentum": 0.9})])
def test_capture_failure_falls_back_to_eager(mode, kw):
    """A CUDA-graph capture that fails (forced here) must not take the round down: the shape is marked
    eager-only and training gives the same result as with graphs switched off."""
    import warnings

    from commefficient_b200.utils import graphs
    w_ref, l_ref, _ = _run("cuda", mode, rounds=5, no_cuda_graphs=True, **kw)
    graphs.FORCE_FAIL = True
    graphs._WARNED.clear()
    try:
        with warnings.catch_warnings(record=True) as caught:
            warnings.simplefilter("always")
            w_f, l_f, _ = _run("cuda", mode, rounds=5, **kw)
    finally:
        graphs.FORCE_FAIL = False
    assert any("capture" in str(c.message) for c in caught), "the fallback must say so once"
    np.testing.assert_allclose(l_f, l_ref, rtol=1e-4, atol=1e-4)
    assert float((w_f - w_ref).norm() / w_ref.norm()) < (5e-2 if mode == "sketch" else 1e-4)
    # and graphs still work afterwards
    w_g, l_g, _ = _run("cuda", mode, rounds=5, **kw)
    np.testing.assert_allclose(l_g, l_ref, rtol=1e-3, atol=1e-3)


def test_failed_capture_with_rng_ops_leaves_the_generator_usable():
    """A capture that dies after random-number kernels were recorded must not poison torch's CUDA generator."""
    from commefficient_b200.utils import graphs

    def body():
        x = torch.randn(1000, device="cuda")          # advances the generator's in-graph offset
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("simulated failure in the middle of a capture")
        return x

    assert graphs.try_capture("cuda", body, what="rng test") is None
    y = torch.randn(1000, device="cuda")              # would raise "Offset increment outside graph capture ..."
    assert torch.isfinite(y).all()
    cap = graphs.try_capture("cuda", lambda: torch.randn(1000, device="cuda"), what="rng test 2")
    assert cap is not None and torch.isfinite(cap.replay()).all()


@pytest.mark.parametrize("mode,kw", [
    ("local_topk", {"error_type": "local", "local_momentum": 0.9}),
    ("fedavg", {"num_fedavg_epochs": 2, "fedavg_batch_size": 3}),
    ("sketch", {"max_grad_norm": 0.3}),              # (small enough that every client is clipped by its own factor)
    ("true_topk", {"do_topk_down": True}),
    ("uncompressed", {"do_dp": True, "dp_mode": "worker", "l2_norm_clip": 1.0, "noise_multiplier": 0.01}),
])
@pytest.mark.parametrize("layout,precision", [("reference", "fp32"), ("auto", "tf32")])
def test_concurrent_client_lanes_equal_sequential_clients(mode, kw, layout, precision):
    """--client_lanes N (clients of a round on N model replicas / streams at once, graphs replayed per lane) gives
    the round the one-after-the-other loop gives: same per-client losses, same weights up to the order of the
    floating-point sum over clients.  Five rounds, so rounds 3-5 run from the per-lane CUDA graphs."""
    seq_w, seq_l, seq_out = _run("cuda", mode, rounds=5, client_lanes=1, flat_layout=layout, precision=precision, **kw)
    par_w, par_l, par_out = _run("cuda", mode, rounds=5, client_lanes=3, flat_layout=layout, precision=precision, **kw)
    np.testing.assert_allclose(par_l, seq_l, rtol=2e-4, atol=1e-5)
    np.testing.assert_allclose(np.asarray(par_out[0]), np.asarray(seq_out[0]), rtol=2e-4, atol=1e-5)
    assert (par_w - seq_w).norm() / seq_w.norm() < 2e-4
