"""Analytic fixtures on the setting of the reference's stale unit test
(/root/reference/CommEfficient/unit_test.py:79-181): linear model y = w.x, sum-of-squares
loss, lr 0.005, X = arange(N*d).view(N, d), y = arange(N).  The first-round values and the
two-parameter k=2 values are the docstring's; its one-parameter second round (0.3808) and
its k=1 second round (-0.3008) do not follow from the update rules of the current reference
code (fed_aggregator.py:511-542,568-613), so those are re-derived here by hand:
  1 param: g(w) = 28(w-1): 0 -> 0.14 -> 0.14 + 0.005*24.08 = 0.2604
  2 params, k=1, virtual error: round 1 picks coord 1 (|-68| > |-56|): w = (0, 0.34),
    Ve = (-56, 0); round 2: g = (-9.76, -10.88), Ve = (-65.76, -10.88) -> coord 0:
    w = (0.3288, 0.34); with virtual momentum 0.9: Vv = (-60.16, -10.88),
    Ve = (-116.16, -10.88) -> w = (0.5808, 0.34)."""
import numpy as np
import pytest
import torch

from commefficient_b200.fed_aggregator import FedModel, FedOptimizer
from commefficient_b200.utils import make_args


class Lin(torch.nn.Module):
    def __init__(self, d):
        super().__init__()
        self.w = torch.nn.Parameter(torch.zeros(d))

    def forward(self, x):
        return x @ self.w


def sum_sq_loss(model, batch, args):
    x, y = batch
    # per-client *mean* loss is what the framework scales by n_i, so mean here = sum / n
    return ((model(x) - y) ** 2).mean(),


def _per_sample(model, batch, args):
    x, y = batch
    return ((model(x) - y) ** 2),


sum_sq_loss.per_sample = _per_sample


# The reference divides the summed transmit by N (=4): sum_i n_i * grad(mean_i) / N = grad(mean over all).
# The fixtures use the SUM loss, i.e. 4x the mean-loss gradient, so lr_eff = 0.005 * 4 on the mean loss.
SCALE = 4.0


@pytest.mark.parametrize("W", [1, 2])
@pytest.mark.parametrize("batching", [True, False])
def test_one_param_uncompressed(W, batching):
    args_lr = 0.005 * SCALE
    N, d = 4, 1
    X = torch.arange(N * d, dtype=torch.float32).view(N, d)
    y = torch.arange(N, dtype=torch.float32)
    ids = torch.arange(N) // (N // W)
    args = make_args(mode="uncompressed", num_workers=W, num_clients=W, local_momentum=0.0, weight_decay=0.0,
                     num_results_train=1, device="cpu", local_batch_size=N // W, no_client_batching=not batching)
    model = Lin(d)
    fm = FedModel(model, sum_sq_loss, args)
    opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args)
    opt.param_groups[0]["lr"] = args_lr
    fm((ids, X, y)); opt.step()
    assert fm.fp.flat.item() == pytest.approx(0.14, abs=1e-6)
    fm((ids, X, y)); opt.step()
    assert fm.fp.flat.item() == pytest.approx(0.2604, abs=1e-6)


def _two_param(W, mode, k, rho=0.0, **kw):
    N, d = 4, 2
    X = torch.arange(N * d, dtype=torch.float32).view(N, d)
    y = torch.arange(N, dtype=torch.float32)
    ids = torch.arange(N) // (N // W)
    args = make_args(mode=mode, k=k, num_rows=9, num_cols=1000, num_blocks=1, num_workers=W, num_clients=W,
                     local_momentum=0.0, virtual_momentum=rho, weight_decay=0.0, error_type="virtual",
                     num_results_train=1, device="cpu", local_batch_size=N // W, **kw)
    model = Lin(d)
    fm = FedModel(model, sum_sq_loss, args)
    opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args)
    opt.param_groups[0]["lr"] = 0.005 * SCALE
    out = []
    for _ in range(2):
        fm((ids, X, y)); opt.step()
        out.append(fm.fp.flat.clone().numpy())
    return out


@pytest.mark.parametrize("W", [1, 2])
@pytest.mark.parametrize("mode", ["sketch", "true_topk"])
def test_two_params_k2(W, mode):
    w1, w2 = _two_param(W, mode, k=2)
    np.testing.assert_allclose(w1, [0.28, 0.34], atol=1e-5)
    np.testing.assert_allclose(w2, [0.172, 0.204], atol=1e-5)


@pytest.mark.parametrize("mode", ["sketch", "true_topk"])
def test_two_workers_k1_error_feedback(mode):
    # k=1: only the larger coordinate moves; the other accumulates in the virtual error
    w1, w2 = _two_param(2, mode, k=1)
    np.testing.assert_allclose(w1, [0.0, 0.34], atol=1e-5)
    np.testing.assert_allclose(w2, [0.3288, 0.34], atol=1e-5)


@pytest.mark.parametrize("mode", ["sketch", "true_topk"])
def test_k1_virtual_momentum(mode):
    w1, w2 = _two_param(2, mode, k=1, rho=0.9)
    np.testing.assert_allclose(w1, [0.0, 0.34], atol=1e-5)
    np.testing.assert_allclose(w2, [0.5808, 0.34], atol=1e-5)
