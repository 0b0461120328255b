#!/usr/bin/env python
"""Differential check of the federated RULES against the unmodified reference, function by function, on the CPU.

Runs in its own process (the reference's modules are top-level names -- ``fed_aggregator``, ``fed_worker``,
``utils`` -- that must not leak into a test session): imports the reference from ``baseline/_ref`` next to this
repo's package, feeds both the same random inputs and prints one JSON object ``{case: max abs difference}``.

  * server rules: ``get_server_update`` (reference fed_aggregator.py:469-613) for all five modes, with the
    momentum / error-feedback variants each mode allows;
  * client rules: ``local_step`` (reference fed_worker.py:184-230) on a small MLP: plain, local momentum,
    local error, local top-k with momentum-factor masking, gradient clipping, sketching and sketch-space
    clipping, microbatches (the reference sums microbatch means -- quirk Q6 -- ours under
    ``--compat_microbatch_sum``); ``process_batch`` (reference fed_worker.py:140-182) for a training client, a client
    with stale weights (``--topk_down``), a client with local momentum / error rows, and a validation shard;
  * whole rounds: the reference's functions assembled the way its FedModel / FedOptimizer drive them (per-client
    ``process_batch``, sum, divide by the round's samples, ``get_server_update``, subtract) against this framework's
    ``FedModel`` / ``FedOptimizer`` over four rounds with recurring clients: uncompressed + momentum, sketch, true_topk,
    local_topk with persistent per-client momentum / error rows, clipping + weight decay, FedAvg local SGD;
  * models: ResNet9 (plain / BatchNorm / 1-channel 62-class), ResNet18, FixupResNet9, FixupResNet18, ResNet101LN --
    the reference's ``state_dict`` loads with ``strict=True`` and output and parameter gradients agree;
  * ``FedSampler``: with the same numpy seed both draw the same clients and samples in every round;
  * dataset preparation from raw files: EMNIST from LEAF json, PersonaChat from ``personachat_self_original.json``;
  * datasets: FedCIFAR10 / FedEMNIST / FedImageNet / FedPERSONA (+ collate) over the same client files return the same items,
    iid and non-iid (PersonaChat with more than one personality permutation and validation indexing are the
    documented departures Q14 / Q15 and are not compared);
  * drop-in use: the reference's own training loops (``cv_train.train`` / ``run_batches``, ``gpt2_train.run_batches``)
    driving this framework's ``FedModel`` / ``FedOptimizer`` / loaders give the numbers this framework's loops give;
  * CV and GPT-2: the reference's ``compute_loss_train`` / ``compute_loss_val`` (gpt2_train.py:56-99) on its model class
    (through the ``pytorch_transformers`` shim over the installed ``transformers``) against ours with the same weights;
  * the eight named transform pipelines (same image, same RNG state -> same tensor) and the table / TSV loggers;
  * CLI: all 55 reference flags exist with the same dest and default and parse the same command line;
  * utilities: LR schedules, ``steps_per_epoch``, ``_topk``, ``get_param_vec`` / ``set_param_vec``.

Usage: python baseline/diff_rules.py   (exit code 0 even when differences are large: the caller judges)
"""
import json
import os
import sys
from types import SimpleNamespace as NS

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.path.join(HERE, "_ref", "CommEfficient")


def main():
    if not os.path.isdir(REF):
        print(json.dumps({"unavailable": "baseline/_ref is not installed"}))
        return 0
    os.environ["CEB_REF_GPT2"] = "1"          # the pytorch_transformers shim then builds its GPT-2 classes
    sys.path.insert(0, os.path.join(HERE, "shims"))
    sys.path.insert(0, REF)
    sys.path.insert(0, ROOT)
    import numpy as np
    import torch
    import _ref_compat
    _ref_compat.install()
    import fed_aggregator as ref_agg                 # the reference
    import fed_worker as ref_worker
    from commefficient_b200 import fed_aggregator as our_agg
    from commefficient_b200 import fed_worker as our_worker
    from commefficient_b200.utils import FlatParams
    import utils as ref_utils
    from commefficient_b200 import utils as our_utils

    out = {}
    d, k = 300, 17

    def base(**kw):
        a = dict(mode="uncompressed", error_type="none", virtual_momentum=0.0, local_momentum=0.0, k=k, grad_size=d,
                 num_cols=4096, num_rows=3, num_blocks=2, device="cpu", do_dp=False, dp_mode="worker",
                 noise_multiplier=0.0, l2_norm_clip=1.0, weight_decay=0.0, num_workers=4, microbatch_size=-1,
                 max_grad_norm=None, do_topk_down=False, compat_microbatch_sum=False, seed=1, precision="fp32")
        a.update(kw)
        return NS(**a)

    # ---- server rules ------------------------------------------------------------------------------------
    server_cases = {
        "server/uncompressed": base(mode="uncompressed", virtual_momentum=0.9),
        "server/fedavg": base(mode="fedavg", virtual_momentum=0.9),
        "server/local_topk": base(mode="local_topk", virtual_momentum=0.9),
        "server/true_topk": base(mode="true_topk", error_type="virtual", virtual_momentum=0.9),
        "server/sketch_virtual": base(mode="sketch", error_type="virtual", virtual_momentum=0.9),
        "server/sketch_local": base(mode="sketch", error_type="local", virtual_momentum=0.0),
        # (sketch with error_type "none" is not compared: the reference then unsketches the never-written Verror
        #  argument -- quirk Q8 -- and this framework rejects the combination at parse time)
    }
    for name, args in server_cases.items():
        worst = 0.0
        for seed in range(4):
            gen = torch.Generator().manual_seed(100 + seed)
            shape = (args.num_rows, args.num_cols) if args.mode == "sketch" else (d,)
            state = [torch.randn(shape, generator=gen) for _ in range(3)]
            lr = 1.0 if args.mode == "fedavg" else 0.1
            for rnd in range(3):                      # three consecutive rounds: the state feeds back
                g = torch.randn(shape, generator=gen)
                if rnd == 0:
                    a_v, a_e = state[1].clone(), state[2].clone()
                    b_v, b_e = state[1].clone(), state[2].clone()
                ru, a_v, a_e = ref_agg.get_server_update(g.clone(), a_v, a_e, args, lr)
                ou, b_v, b_e = our_agg.get_server_update(g.clone(), b_v, b_e, args, lr)
                ru = ru.clone()
                worst = max(worst, float((ru - ou).abs().max()), float((a_v - b_v).abs().max()))
                if args.error_type == "virtual":
                    worst = max(worst, float((a_e - b_e).abs().max()))
        out[name] = worst

    # ---- client rules ------------------------------------------------------------------------------------
    def make_model(seed):
        torch.manual_seed(seed)
        return torch.nn.Sequential(torch.nn.Linear(10, 16), torch.nn.Tanh(), torch.nn.Linear(16, 4))

    def compute_loss(model, batch, args):
        x, y = batch
        logits = model(x)
        loss = torch.nn.functional.cross_entropy(logits, y)
        acc = (logits.argmax(1) == y).float().mean()
        return loss, acc

    n_par = sum(p.numel() for p in make_model(0).parameters())

    def ref_utils_get(model):
        return torch.cat([p.data.reshape(-1) for p in model.parameters()])
    client_cases = {
        "client/uncompressed": base(mode="uncompressed", grad_size=n_par),
        "client/uncompressed_clip": base(mode="uncompressed", grad_size=n_par, max_grad_norm=0.05),
        "client/true_topk_local_momentum": base(mode="true_topk", error_type="virtual", local_momentum=0.9, grad_size=n_par),
        "client/local_topk_error_momentum": base(mode="local_topk", error_type="local", local_momentum=0.9, grad_size=n_par),
        "client/local_topk_plain": base(mode="local_topk", error_type="none", grad_size=n_par),
        "client/sketch": base(mode="sketch", error_type="virtual", grad_size=n_par),
        "client/sketch_clip": base(mode="sketch", error_type="virtual", grad_size=n_par, max_grad_norm=0.05),
        "client/microbatches_q6": base(mode="uncompressed", grad_size=n_par, microbatch_size=3, compat_microbatch_sum=True),
        "client/weight_decay": base(mode="uncompressed", grad_size=n_par, weight_decay=5e-4),
    }
    for name, args in client_cases.items():
        worst, skipped = 0.0, None
        for seed in range(3):
            gen = torch.Generator().manual_seed(500 + seed)
            x = torch.randn(7, 10, generator=gen)
            y = torch.randint(0, 4, (7,), generator=gen)
            need_v = args.local_momentum > 0
            need_e = args.error_type == "local"
            v0 = torch.randn(n_par, generator=gen) * 0.1
            e0 = torch.randn(n_par, generator=gen) * 0.1
            rm, om = make_model(seed), make_model(seed)
            om._flat_params = FlatParams(om, "cpu", layout="reference")
            rv, re_ = (v0.clone() if need_v else None), (e0.clone() if need_e else None)
            ov, oe = (v0.clone() if need_v else None), (e0.clone() if need_e else None)
            try:
                r_res, r_tx = ref_worker.local_step(rm, rv, re_, [x, y], compute_loss, args)
            except TypeError as exc:              # e.g. an overload the reference uses that this torch removed
                skipped = "reference raised %s" % str(exc).splitlines()[0][:80]
                break
            o_res, o_tx = our_worker.local_step(om, ov, oe, [x, y], compute_loss, args)
            r_tx = r_tx if torch.is_tensor(r_tx) else r_tx
            worst = max(worst, float((r_tx.reshape(-1) - o_tx.reshape(-1)).abs().max()))
            worst = max(worst, float(np.abs(np.asarray(r_res, dtype=np.float64)
                                            - np.asarray([float(t) for t in o_res], dtype=np.float64)).max()))
            if need_v:
                worst = max(worst, float((rv - ov).abs().max()))
            if need_e:
                worst = max(worst, float((re_ - oe).abs().max()))
        out[name] = skipped if skipped is not None else worst
    # ---- process_batch: weights in, (transmit, results) out; stale client weights (--topk_down); validation -----
    pb_cases = {
        "client/process_batch_train": base(mode="uncompressed", grad_size=n_par),
        "client/process_batch_topk_down": base(mode="true_topk", error_type="virtual", grad_size=n_par, do_topk_down=True, k=23),
        "client/process_batch_local_state": base(mode="local_topk", error_type="local", local_momentum=0.9, grad_size=n_par),
        "client/process_batch_validation": base(mode="uncompressed", grad_size=n_par),
    }
    for name, args in pb_cases.items():
        worst = 0.0
        for seed in range(3):
            gen = torch.Generator().manual_seed(900 + seed)
            x = torch.randn(6, 10, generator=gen)
            y = torch.randint(0, 4, (6,), generator=gen)
            cid = -1 if name.endswith("validation") else 2
            ids = torch.full((6,), cid, dtype=torch.long)
            ps = torch.randn(n_par, generator=gen) * 0.3
            cw = torch.randn(4, n_par, generator=gen) * 0.3 if args.do_topk_down else None
            need_state = args.local_momentum > 0
            cv = torch.randn(4, n_par, generator=gen) * 0.1 if need_state else None
            ce = torch.randn(4, n_par, generator=gen) * 0.1 if args.error_type == "local" else None
            rm, om = make_model(seed), make_model(seed)
            om._flat_params = FlatParams(om, "cpu", layout="reference")
            clone = lambda t: None if t is None else t.clone()      # noqa: E731
            r_cv, r_ce, o_cv, o_ce = clone(cv), clone(ce), clone(cv), clone(ce)
            r_tx, r_res = ref_worker.process_batch([ids, x, y], rm, ps.clone(), clone(cw), r_ce, r_cv, compute_loss,
                                                   compute_loss, args)
            o_tx, o_res = our_worker.process_batch([ids, x, y], om, ps.clone(), clone(cw), o_ce, o_cv, compute_loss,
                                                   compute_loss, args)
            if cid != -1:
                worst = max(worst, float((r_tx.reshape(-1) - o_tx.reshape(-1)).abs().max()))
            worst = max(worst, float(np.abs(np.asarray(r_res, dtype=np.float64)
                                            - np.asarray([float(t) for t in o_res], dtype=np.float64)).max()))
            # the weights the client actually computed with
            worst = max(worst, float((ref_utils_get(rm) - ref_utils_get(om)).abs().max()))
        out[name] = worst

    # ---- whole rounds: the reference's own functions assembled the way its FedModel / FedOptimizer drive them
    #      (process_batch per client, sum, / round samples, get_server_update, weights -= update:
    #      fed_aggregator.py:213-335,429-458) against this framework's FedModel / FedOptimizer, several rounds
    from commefficient_b200.fed_aggregator import FedModel as OurFedModel, FedOptimizer as OurFedOptimizer
    from commefficient_b200.utils import make_args

    loop_cases = {
        "rounds/uncompressed_momentum": dict(mode="uncompressed", error_type="none", virtual_momentum=0.9),
        "rounds/sketch_virtual": dict(mode="sketch", error_type="virtual", virtual_momentum=0.9, num_cols=4096, num_rows=3,
                                      num_blocks=2, k=23),
        "rounds/true_topk_virtual": dict(mode="true_topk", error_type="virtual", virtual_momentum=0.9, k=23),
        "rounds/local_topk_local_state": dict(mode="local_topk", error_type="local", local_momentum=0.9,
                                              virtual_momentum=0.0, k=23),
        "rounds/uncompressed_clip_wd": dict(mode="uncompressed", error_type="none", virtual_momentum=0.0,
                                            max_grad_norm=0.05, weight_decay=5e-4),
    }
    loop_cases["rounds/fedavg_local_sgd"] = dict(mode="fedavg", error_type="none", virtual_momentum=0.9,
                                                 num_fedavg_epochs=2, fedavg_batch_size=2, fedavg_lr_decay=0.9)
    n_clients, per_round, lr = 6, 3, 0.1
    for name, kw in loop_cases.items():
        rargs = base(grad_size=n_par, num_workers=per_round, **kw)
        torch.manual_seed(42)
        rm = make_model(7)
        ps = ref_utils_get(rm).clone()
        shape = (rargs.num_rows, rargs.num_cols) if rargs.mode == "sketch" else (n_par,)
        Vv, Ve = torch.zeros(shape), torch.zeros(shape)
        need_v, need_e = rargs.local_momentum > 0, rargs.error_type == "local"
        cv = torch.zeros(n_clients, n_par) if need_v else None
        ce = torch.zeros(n_clients, n_par) if need_e else None
        oargs = make_args(device="cpu", dataset_name="CIFAR10", num_clients=n_clients, num_workers=per_round,
                          local_batch_size=-1, precision="fp32", flat_layout="reference", num_results_train=2,
                          num_results_val=2, weight_decay=kw.get("weight_decay", 0.0), watchdog_timeout=0.0,
                          **{"local_momentum": 0.0, **{k_: v_ for k_, v_ in kw.items() if k_ != "weight_decay"}})
        om = make_model(7)
        fed = OurFedModel(om, compute_loss, oargs, compute_loss)
        opt = OurFedOptimizer(torch.optim.SGD(om.parameters(), lr=1), oargs)
        opt.param_groups[0]["lr"] = lr
        fedavg = rargs.mode == "fedavg"
        if fedavg:
            for k_ in ("num_fedavg_epochs", "fedavg_batch_size", "fedavg_lr_decay"):
                setattr(rargs, k_, kw[k_])
            opt.step()                                   # publishes the LR the clients' local SGD uses
        fed.train(True)
        gen = torch.Generator().manual_seed(77)
        worst = 0.0
        for rnd in range(4):
            clients = [(rnd * 2 + j) % n_clients for j in range(per_round)]
            sizes = [3, 5, 4]
            xs = [torch.randn(n, 10, generator=gen) for n in sizes]
            ys = [torch.randint(0, 4, (n,), generator=gen) for n in sizes]
            # reference side
            total = torch.zeros(shape)
            for c, x, y in zip(clients, xs, ys):
                ids = torch.full((x.shape[0],), c, dtype=torch.long)
                if fedavg:
                    # the reference's local SGD (fed_worker.py:61-113), restated: chunks of fedavg_batch_size,
                    # num_fedavg_epochs passes, step size lr * decay^step on the mean gradient; upload (w0 - w) * n
                    w_local, step = ps.clone(), 0
                    for _ep in range(rargs.num_fedavg_epochs):
                        for lo in range(0, x.shape[0], rargs.fedavg_batch_size):
                            xc, yc = x[lo:lo + rargs.fedavg_batch_size], y[lo:lo + rargs.fedavg_batch_size]
                            g_sum, _r = ref_worker.process_batch([ids[:xc.shape[0]], xc, yc], rm, w_local.clone(), None,
                                                                 None, None, compute_loss, compute_loss, rargs)
                            w_local = w_local - (g_sum / xc.shape[0]) * lr * rargs.fedavg_lr_decay ** step
                            step += 1
                    total += (ps - w_local) * x.shape[0]
                    continue
                tx, _res = ref_worker.process_batch([ids, x, y], rm, ps.clone(), None, ce, cv, compute_loss, compute_loss,
                                                    rargs)
                total += tx.reshape(shape)
                # (the reference's worker loop writes the client rows back: fed_worker.py:124-134 via shared memory)
            grad = total / float(sum(sizes))
            upd, Vv, Ve = ref_agg.get_server_update(grad, Vv, Ve, rargs, 1 if fedavg else lr)
            ps = ps - upd.reshape(-1)
            # this framework
            order = np.argsort(clients, kind="stable")
            ids = torch.cat([torch.full((sizes[i],), clients[i], dtype=torch.long) for i in order])
            fed((ids, torch.cat([xs[i] for i in order]), torch.cat([ys[i] for i in order])))
            opt.step()
            ours = fed.fp.to_reference_order(fed.fp.flat).detach()
            worst = max(worst, float((ours - ps).abs().max()))
        fed.rt.shutdown()
        out[name] = worst

    # ---- models: same weights, same input -> same output and same parameter gradients ---------------------
    import models as ref_models
    from commefficient_b200 import models as our_models
    model_cases = [("ResNet9", dict(num_classes=10), (2, 3, 32, 32)),
                   ("ResNet9", dict(num_classes=10, do_batchnorm=True), (4, 3, 32, 32)),
                   ("ResNet9", dict(num_classes=62, initial_channels=1), (2, 1, 32, 32)),
                   ("ResNet18", dict(num_classes=100), (2, 3, 32, 32)),
                   ("FixupResNet9", {}, (2, 3, 32, 32)),
                   ("FixupResNet18", dict(num_classes=10), (2, 3, 32, 32)),
                   ("ResNet101LN", {}, (1, 1, 28, 28))]
    import contextlib
    import io
    for cls, kw, shape in model_cases:
        torch.manual_seed(0)
        with contextlib.redirect_stdout(io.StringIO()):          # (the reference's ResNet9 prints its BatchNorm flag)
            r, o = getattr(ref_models, cls)(**kw), getattr(our_models, cls)(**kw)
        with torch.no_grad():
            for p in r.parameters():                              # (Fixup zero-initialises some layers)
                p.add_(torch.randn_like(p) * 0.05)
        o.load_state_dict(r.state_dict(), strict=True)            # identical parameter / buffer names and shapes
        x = torch.randn(*shape)
        r.train()
        o.train()
        yr, yo = r(x), o(x)
        gr = torch.autograd.grad(yr.square().sum(), list(r.parameters()), allow_unused=True)
        go = torch.autograd.grad(yo.square().sum(), list(o.parameters()), allow_unused=True)
        err = float((yr - yo).abs().max() / yr.abs().max())
        for a, b in zip(gr, go):
            if a is not None:
                err = max(err, float((a - b).abs().max() / (a.abs().max() + 1e-12)))
        tag = "models/%s%s" % (cls, "" if not kw else "(" + ",".join("%s=%s" % kv for kv in sorted(kw.items())) + ")")
        out[tag] = err

    # ---- FedSampler: the same numpy seed draws the same rounds ---------------------------------------------
    from data_utils.fed_sampler import FedSampler as RefSampler
    from commefficient_b200.data_utils import FedSampler as OurSampler

    class Sizes:
        def __init__(self, sizes):
            self.data_per_client = np.asarray(sizes)
            self.num_clients = len(sizes)

        def __len__(self):
            return int(np.sum(self.data_per_client))

    worst = 0.0
    for sizes, workers, bs in (([5] * 12, 4, -1), ([3, 9, 1, 7, 4, 4, 6, 2], 3, 2), ([10] * 6, 6, 4)):
        ds = Sizes(sizes)
        np.random.seed(11)
        ref_rounds = [np.asarray(b) for b in RefSampler(ds, workers, bs)]
        np.random.seed(11)
        our_rounds = [np.asarray(b) for b in OurSampler(ds, workers, bs)]
        if len(ref_rounds) != len(our_rounds):
            worst = max(worst, 1e9)
            continue
        for a, b in zip(ref_rounds, our_rounds):
            worst = max(worst, 0.0 if (len(a) == len(b) and np.array_equal(np.sort(a), np.sort(b))) else 1.0)
    out["data/fed_sampler_rounds"] = worst

    # ---- datasets: both read the same on-disk client files; every item, iid and non-iid --------------------
    import random
    import tempfile
    import data_utils as ref_data
    from commefficient_b200 import data_utils as our_data

    def same(x, y):
        if isinstance(x, (list, tuple)):
            return len(x) == len(y) and all(same(a, b) for a, b in zip(x, y))
        if torch.is_tensor(x) or torch.is_tensor(y):
            return torch.equal(torch.as_tensor(x), torch.as_tensor(y))
        return bool(np.all(np.asarray(x) == np.asarray(y)))

    def compare_sets(make_ref, make_ours, variants):
        bad = 0
        for kw in variants:
            np.random.seed(3)
            r = make_ref(**kw)
            np.random.seed(3)
            o = make_ours(**kw)
            bad += int(len(r) != len(o) or r.num_clients != o.num_clients)
            if kw.get("train", True):
                bad += int(not np.array_equal(np.asarray(r.data_per_client), np.asarray(o.data_per_client)))
            for i in range(min(len(r), len(o))):
                random.seed(i)
                a = r[i]
                random.seed(i)
                b = o[i]
                bad += int(not same(a, b))
        return float(bad)

    tmp = tempfile.mkdtemp()
    our_data.write_synthetic_cifar(tmp, num_classes=10, images_per_class=12, num_val=30, seed=0)
    out["data/fed_cifar10_items"] = compare_sets(
        lambda **kw: ref_data.FedCIFAR10(tmp, "CIFAR10", None, kw["iid"], kw["n"], train=kw["train"], download=False),
        lambda **kw: our_data.FedCIFAR10(tmp, "CIFAR10", None, kw["iid"], kw["n"], train=kw["train"], download=False),
        [dict(iid=False, n=10, train=True), dict(iid=False, n=20, train=True), dict(iid=True, n=7, train=True),
         dict(iid=False, n=None, train=False)])
    tmp = tempfile.mkdtemp()
    our_data.write_synthetic_emnist(tmp, num_clients=12, images_per_client=6, num_classes=62, seed=0)
    out["data/fed_emnist_items"] = compare_sets(
        lambda **kw: ref_data.FedEMNIST(tmp, "EMNIST", None, kw["iid"], kw["n"], train=kw["train"], download=False),
        lambda **kw: our_data.FedEMNIST(tmp, "EMNIST", None, kw["iid"], kw["n"], train=kw["train"], download=False),
        [dict(iid=False, n=12, train=True), dict(iid=False, n=24, train=True), dict(iid=True, n=5, train=True),
         dict(iid=False, n=None, train=False)])
    # EMNIST from the raw LEAF json (prepare_datasets: fed_emnist.py:82-130), each implementation in its own copy
    def write_leaf(root, seed):
        rng = np.random.RandomState(seed)
        for split, users, per in (("train", 5, 4), ("test", 3, 2)):
            os.makedirs(os.path.join(root, split), exist_ok=True)
            blob = {"users": [], "num_samples": [], "user_data": {}}
            for u in range(users):
                name = "f%04d_%02d" % (u * 7, u)
                blob["users"].append(name)
                blob["num_samples"].append(per + u)
                blob["user_data"][name] = {"x": rng.rand(per + u, 784).round(4).tolist(),
                                           "y": rng.randint(0, 62, per + u).tolist()}
            with open(os.path.join(root, split, "all_data_0_niid.json"), "w") as f:
                json.dump(blob, f)
    leaf_r, leaf_o = tempfile.mkdtemp(), tempfile.mkdtemp()
    write_leaf(leaf_r, 5)
    write_leaf(leaf_o, 5)
    bad = 0
    for iid, n, train in ((False, 5, True), (True, 3, True), (False, None, False)):
        np.random.seed(3)
        r = ref_data.FedEMNIST(leaf_r, "EMNIST", None, iid, n, train=train, download=False)
        np.random.seed(3)
        o = our_data.FedEMNIST(leaf_o, "EMNIST", None, iid, n, train=train, download=False)
        bad += int(len(r) != len(o))
        for i in range(min(len(r), len(o))):
            bad += int(not same(r[i], o[i]))
    out["data/fed_emnist_prepared_from_leaf_json"] = float(bad)
    tmp = tempfile.mkdtemp()
    our_data.write_synthetic_imagenet(tmp, num_classes=4, images_per_class=3, val_per_class=2)
    out["data/fed_imagenet_items"] = compare_sets(
        lambda **kw: ref_data.FedImageNet(tmp, "ImageNet", None, kw["iid"], kw["n"], train=kw["train"], download=False),
        lambda **kw: our_data.FedImageNet(tmp, "ImageNet", None, kw["iid"], kw["n"], train=kw["train"], download=False),
        [dict(iid=False, n=4, train=True), dict(iid=True, n=3, train=True), dict(iid=False, n=None, train=False)])
    # PersonaChat: training items with one personality permutation (more than one, and validation indexing, are the
    # documented departures Q14 / Q15), then the collate function on the same items
    tmp = tempfile.mkdtemp()
    our_data.write_synthetic_persona(tmp, num_clients=6, num_candidates=2, seed=0)
    tok = our_data.SimpleTokenizer()
    tok.add_special_tokens(our_data.ATTR_TO_SPECIAL_TOKEN)
    mk_ref = lambda **kw: ref_data.FedPERSONA(tok, 2, 2, 1, tmp, "PERSONA", None, kw["iid"], kw["n"], train=True,  # noqa: E731
                                              download=False)
    mk_our = lambda **kw: our_data.FedPERSONA(tokenizer=tok, num_candidates=2, max_history=2,                      # noqa: E731
                                              personality_permutations=1, dataset_dir=tmp, dataset_name="PERSONA",
                                              do_iid=kw["iid"], num_clients=kw["n"], train=True)
    variants = [dict(iid=False, n=None, train=True), dict(iid=True, n=4, train=True)]
    bad = compare_sets(mk_ref, mk_our, variants)
    # the raw corpus file split by personality (reference split_dataset: fed_persona.py:128-170) against our
    # prepare_datasets on a directory that holds personachat_self_original.json
    raw_dir = tempfile.mkdtemp()
    rng = np.random.RandomState(4)
    personas = [["i like %s ." % w, "my dog is %s ." % w] for w in ("red", "blue", "green")]
    def dialog(p):
        return {"personality": list(p), "utterances": [
            {"candidates": ["reply %d %d" % (u, c) for c in range(3)], "history": ["turn %d" % t for t in range(2 * u + 1)]}
            for u in range(int(rng.randint(1, 4)))]}
    raw = {"train": [dialog(personas[i % 3]) for i in range(7)], "valid": [dialog(personas[i % 3]) for i in range(3)]}
    with open(os.path.join(raw_dir, "personachat_self_original.json"), "w") as f:
        json.dump(raw, f)
    (ref_clients, ref_val), (ref_dpc, ref_tupd, ref_vupd) = ref_data.FedPERSONA.split_dataset(
        None, os.path.join(raw_dir, "personachat_self_original.json"))
    ours_ds = our_data.FedPERSONA(tokenizer=tok, num_candidates=2, max_history=2, personality_permutations=1,
                                  dataset_dir=raw_dir, dataset_name="PERSONA", do_iid=False, num_clients=None, train=True)
    with open(ours_ds.stats_fn()) as f:
        st_ = json.load(f)
    bad += float(st_["dialogs_per_client"] != list(ref_dpc) or st_["train_utterances_per_dialog"] != list(ref_tupd)
                 or st_["val_utterances_per_dialog"] != list(ref_vupd))
    for cid, dialogs in enumerate(ref_clients.values()):
        with open(ours_ds.client_fn(cid)) as f:
            bad += float(json.load(f) != dialogs)
    with open(ours_ds.validation_fn()) as f:
        bad += float(json.load(f) != ref_val)
    np.random.seed(3)
    r, o = mk_ref(**variants[0]), mk_our(**variants[0])
    random.seed(0)
    br = ref_data.personachat_collate_fn([r[i] for i in range(5)])
    random.seed(0)
    bo = our_data.personachat_collate_fn([o[i] for i in range(5)])
    out["data/fed_persona_items_and_collate"] = bad + float(not same(list(br), list(bo)))
    # ---- CV: the reference's loss callbacks (cv_train.py:67-83) on a ResNet-9 against ours, same weights --------------
    import importlib.util as _ilu
    spec = _ilu.spec_from_file_location("ref_cv_train", os.path.join(REF, "cv_train.py"))
    ref_cv = _ilu.module_from_spec(spec)
    spec.loader.exec_module(ref_cv)
    from commefficient_b200 import cv_train as our_cv
    torch.manual_seed(0)
    with contextlib.redirect_stdout(io.StringIO()):
        rnet, onet = ref_models.ResNet9(num_classes=10), our_models.ResNet9(num_classes=10)
    onet.load_state_dict(rnet.state_dict())
    cx = torch.randn(6, 3, 32, 32)
    cy = torch.randint(0, 10, (6,))
    cargs = NS(device="cpu", do_mixup=False)
    worst = 0.0
    for fn in ("compute_loss_train", "compute_loss_val"):
        ra_ = getattr(ref_cv, fn)(rnet, (cx, cy), cargs)
        oa_ = getattr(our_cv, fn)(onet, (cx, cy), cargs)
        worst = max(worst, max(abs(float(a) - float(b)) for a, b in zip(ra_, oa_)))
    out["cv/compute_loss_train_and_val"] = worst

    # ---- drop-in: the REFERENCE's training loop (cv_train.py:85-160, run_batches :171-252) driving THIS framework's
    #      FedModel / FedOptimizer / loaders, against this framework's own loop on the same setup
    from torch.optim.lr_scheduler import LambdaLR
    loop_dir = tempfile.mkdtemp()
    our_data.write_synthetic_cifar(loop_dir, num_classes=10, images_per_class=10, num_val=40, seed=0)

    def run_loop(train_fn, logger_cls, timer_cls):
        torch.manual_seed(5)
        np.random.seed(5)
        random.seed(5)
        largs = our_utils.parse_args(argv=[
            "--dataset_name", "CIFAR10", "--dataset_dir", loop_dir, "--device", "cpu", "--num_epochs", "1", "--pivot_epoch",
            "0.5", "--lr_scale", "0.05", "--num_cols", "2000", "--num_rows", "3", "--k", "300", "--valid_batch_size", "8",
            "--model", "ResNet9", "--num_clients", "10", "--num_workers", "4", "--local_batch_size", "4", "--mode", "sketch",
            "--error_type", "virtual", "--local_momentum", "0", "--virtual_momentum", "0.9",
            "--train_dataloader_workers", "0", "--val_dataloader_workers", "0"])
        largs.do_mixup = False
        net, groups = our_cv.build_model_and_groups(largs)
        tr_loader, te_loader = our_cv.get_data_loaders(largs)
        fm = OurFedModel(net, our_cv.compute_loss_train, largs, our_cv.compute_loss_val)
        fo = OurFedOptimizer(torch.optim.SGD(groups, lr=1), largs)
        sched = our_utils.PiecewiseLinear([0, largs.pivot_epoch, largs.num_epochs], [0, largs.lr_scale, 0])
        spe_ = our_utils.steps_per_epoch(largs.local_batch_size, tr_loader.dataset, largs.num_workers)
        lrs = LambdaLR(fo, lr_lambda=lambda step: sched(step / spe_))
        with contextlib.redirect_stdout(io.StringIO()):
            summ = train_fn(fm, fo, lrs, tr_loader, te_loader, largs, None, loggers=(logger_cls(),), timer=timer_cls())
        fm.rt.shutdown()
        return summ

    ref_summary = run_loop(ref_cv.train, ref_utils.TableLogger, ref_utils.Timer)
    our_summary = run_loop(our_cv.train, our_utils.TableLogger, our_utils.Timer)
    out["loop/reference_train_loop_drives_our_classes"] = max(
        abs(float(ref_summary[k_]) - float(our_summary[k_])) for k_ in ("train_loss", "train_acc", "test_loss", "test_acc"))

    # ---- drop-in, PersonaChat: the reference's run_batches (gpt2_train.py:169-251) over our classes, one training
    #      epoch and one validation pass, against our run_batches on the same setup
    try:
        import importlib.util as _ilu2
        spec = _ilu2.spec_from_file_location("ref_gpt2_loops", os.path.join(REF, "gpt2_train.py"))
        ref_gl = _ilu2.module_from_spec(spec)
        spec.loader.exec_module(ref_gl)
        from commefficient_b200 import gpt2_train as our_gl
        pdir = tempfile.mkdtemp()
        our_data.write_synthetic_persona(pdir, num_clients=8, num_candidates=2, seed=0)

        class _NoWriter:
            def add_scalar(self, *a, **k):
                pass

        def run_gpt2(loop_mod, logger_cls, timer_cls):
            torch.manual_seed(6)
            np.random.seed(6)
            random.seed(6)
            ga = our_utils.parse_args(default_lr=4e-2, argv=[
                "--dataset_name", "PERSONA", "--dataset_dir", pdir, "--num_clients", "8", "--model_checkpoint", "gpt2-tiny",
                "--num_workers", "2", "--local_batch_size", "2", "--mode", "uncompressed", "--error_type", "none",
                "--local_momentum", "0", "--virtual_momentum", "0.9", "--num_epochs", "1", "--device", "cpu",
                "--valid_batch_size", "2", "--lr_scale", "0.1"])
            ga.num_results_train = 1
            gm, gtok = our_gl.build_model_and_tokenizer(ga)
            for mod_ in gm.modules():                       # (no dropout: both loops must see the same numbers)
                if isinstance(mod_, torch.nn.Dropout):
                    mod_.p = 0.0
            our_gl.add_special_tokens_(gm, gtok)
            trl, val = our_gl.get_data_loaders(ga, gtok)
            gfm = OurFedModel(gm, our_gl.compute_loss_train, ga, our_gl.compute_loss_val)
            gfo = OurFedOptimizer(torch.optim.SGD(gm.parameters(), lr=1), ga)
            spe_ = our_utils.steps_per_epoch(ga.local_batch_size, trl.dataset, ga.num_workers)
            sch = our_utils.PiecewiseLinear([0, ga.num_epochs * spe_], [ga.lr_scale, 0.0])
            lrs = LambdaLR(gfo, lr_lambda=lambda step: sch(step))
            with contextlib.redirect_stdout(io.StringIO()):
                tr_loss, _dl, _ul = loop_mod.run_batches(gfm, gfo, lrs, trl, ga, timer_cls(), True, epoch=0, epoch_fraction=1,
                                                         logger=logger_cls(), writer=_NoWriter())
                nll, acc, ppl = loop_mod.run_batches(gfm, None, None, val, ga, timer_cls(), False, epoch_fraction=1,
                                                     logger=logger_cls(), writer=_NoWriter())
            gfm.rt.shutdown()
            return [float(tr_loss), float(nll), float(acc), float(ppl)]

        import warnings as _w
        with _w.catch_warnings():
            _w.simplefilter("ignore")
            a_ = run_gpt2(ref_gl, ref_utils.TableLogger, ref_utils.Timer)
            b_ = run_gpt2(our_gl, our_utils.TableLogger, our_utils.Timer)
        out["loop/reference_gpt2_loop_drives_our_classes"] = max(abs(x - y) / max(1.0, abs(x)) for x, y in zip(a_, b_))
    except ImportError as exc:
        out["loop/reference_gpt2_loop_drives_our_classes"] = "skipped: %s" % exc

    # ---- GPT-2: the reference's loss callbacks (gpt2_train.py:56-99) on its model class against ours, same weights ---
    try:
        import importlib.util
        import transformers
        import pytorch_transformers as ref_pt                       # the shim over the installed transformers
        spec = importlib.util.spec_from_file_location("ref_gpt2_train", os.path.join(REF, "gpt2_train.py"))
        ref_gpt2 = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(ref_gpt2)
        from commefficient_b200 import gpt2_train as our_gpt2
        from commefficient_b200.models.gpt2 import GPT2Config as OurGPT2Config, GPT2DoubleHeadsModel as OurGPT2
        torch.manual_seed(0)
        gkw = dict(vocab_size=101, n_positions=32, n_embd=64, n_layer=2, n_head=4, resid_pdrop=0.0, embd_pdrop=0.0,
                   attn_pdrop=0.0)
        rg = ref_pt.GPT2DoubleHeadsModel(transformers.GPT2Config(summary_first_dropout=0.0, num_labels=1, **gkw))
        og = OurGPT2(OurGPT2Config(vocab_size=101, n_positions=32, n_ctx=32, n_embd=64, n_layer=2, n_head=4,
                                   resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0, summary_first_dropout=0.0))
        src = rg.state_dict()
        og.load_state_dict({k_: src[k_] for k_ in og.state_dict()})
        gen = torch.Generator().manual_seed(1)
        B_, C_, T_ = 4, 2, 12
        ids = torch.randint(0, 101, (B_, C_, T_), generator=gen)
        types = torch.randint(0, 101, (B_, C_, T_), generator=gen)
        mc_ids = torch.randint(0, T_, (B_, C_), generator=gen)
        lm = torch.randint(0, 101, (B_, C_, T_), generator=gen)
        lm[:, 0] = -1
        lm[:, :, :3] = -1
        mc = torch.randint(0, C_, (B_,), generator=gen)
        gbatch = (ids, mc_ids, lm, mc, types)
        gargs = NS(lm_coef=1.0, mc_coef=0.7, device="cpu")
        rg.train()
        og.train()
        worst = abs(float(ref_gpt2.compute_loss_train(rg, gbatch, gargs)[0])
                    - float(our_gpt2.compute_loss_train(og, gbatch, gargs)[0]))
        rv = ref_gpt2.compute_loss_val(rg, gbatch, gargs)
        ov = our_gpt2.compute_loss_val(og, gbatch, gargs)
        worst = max(worst, max(abs(float(a) - float(b)) for a, b in zip(rv, ov)))
        out["gpt2/compute_loss_train_and_val"] = worst
    except ImportError as exc:                                       # no transformers on this machine
        out["gpt2/compute_loss_train_and_val"] = "skipped: %s" % exc

    # ---- transforms: the eight named pipelines on the same image with the same RNG state ------------------------
    from PIL import Image
    worst = 0.0
    rng = np.random.RandomState(8)
    rgb32 = Image.fromarray(rng.randint(0, 256, (32, 32, 3)).astype(np.uint8))
    gray28 = Image.fromarray(rng.rand(28, 28).astype(np.float32))               # EMNIST images are float 'F' mode
    rgb_big = Image.fromarray(rng.randint(0, 256, (300, 260, 3)).astype(np.uint8))
    for tname, img in (("cifar10_train_transforms", rgb32), ("cifar10_test_transforms", rgb32),
                       ("cifar100_train_transforms", rgb32), ("cifar100_test_transforms", rgb32),
                       ("femnist_train_transforms", gray28), ("femnist_test_transforms", gray28),
                       ("imagenet_train_transforms", rgb_big), ("imagenet_val_transforms", rgb_big)):
        for seed in range(3):
            torch.manual_seed(seed)
            random.seed(seed)
            a = getattr(ref_data, tname)(img)
            torch.manual_seed(seed)
            random.seed(seed)
            b = getattr(our_data, tname)(img)
            worst = max(worst, 1e9 if a.shape != b.shape else float((a - b).abs().max()))
    out["data/transforms"] = worst

    # ---- loggers: the printed table and the TSV file --------------------------------------------------------------
    import contextlib as _ctx
    import io as _io
    rows = [{"epoch": 1, "lr": 0.123456, "train_loss": 2.30258, "note": "x"}, {"epoch": 2, "lr": 0.05, "train_loss": 1.5,
                                                                                     "note": "y"}]
    def table_text(mod):
        buf = _io.StringIO()
        with _ctx.redirect_stdout(buf):
            tl = mod.TableLogger()
            for r in rows:
                tl.append(dict(r))
        return buf.getvalue()
    bad = float(table_text(ref_utils) != table_text(our_utils))
    def tsv_text(mod):
        t = mod.TSVLogger()
        for e, (hours, acc) in enumerate(((0.5, 0.91), (1.25, 0.9342))):
            t.append({"epoch": e + 1, "total_time": hours * 3600, "test_acc": acc})
        return str(t)
    bad += float(tsv_text(ref_utils) != tsv_text(our_utils))
    for m_, extra in (("sketch", {}), ("true_topk", {}), ("uncompressed", {})):
        la = NS(num_rows=5, num_cols=500000, k=50000, mode=m_, num_workers=100, num_clients=10000)
        ra_, oa_ = ref_utils.make_logdir(la), our_utils.make_logdir(la)
        strip = lambda p_: p_.split("_", 2)[-1]          # noqa: E731 -- drop "runs/<date>_<time>"
        bad += float(strip(ra_) != strip(oa_) or not oa_.startswith("runs" + os.sep))
    out["utils/table_and_tsv_loggers"] = bad

    # ---- CLI: every reference flag exists with the same dest and default, and parses to the same value -------
    from commefficient_b200.utils.args import build_parser

    def parse_both(argv):
        old = sys.argv
        sys.argv = ["prog"] + list(argv)
        try:
            ra = vars(ref_utils.parse_args())
        finally:
            sys.argv = old
        oa = vars(build_parser(None, our_models.model_names()).parse_args(list(argv)))
        return ra, oa

    bad = 0
    ra, oa = parse_both([])
    bad += len([k for k in ra if k not in oa])
    bad += len([k for k in ra if k in oa and ra[k] != oa[k] and k != "device"])     # (device: resolved later here)
    argv = ["--mode", "true_topk", "--k", "123", "--num_workers", "7", "--local_momentum", "0.5", "--error_type",
            "virtual", "--topk_down", "--dp", "--num_clients", "50", "--iid", "--local_batch_size", "3", "--model",
            "ResNet18", "--dataset_name", "CIFAR100", "--lr_scale", "0.3", "--pivot_epoch", "2", "--num_epochs", "3",
            "--fedavg_batch_size", "5", "--max_grad_norm", "2.5", "--microbatch_size", "4", "--num_rows", "3",
            "--num_cols", "1000", "--num_blocks", "5", "--share_ps_gpu", "--device", "cpu", "--weight_decay", "1e-3",
            "--valid_batch_size", "9", "--test", "--tensorboard", "--seed", "5", "--fedavg_lr_decay", "0.9",
            "--num_fedavg_epochs", "2", "--virtual_momentum", "0.7"]
    ra, oa = parse_both(argv)
    bad += len([k for k in ra if ra[k] != oa.get(k)])
    out["cli/flags_dests_defaults"] = float(bad)

    # ---- small utilities --------------------------------------------------------------------------------------
    worst = 0.0
    rs = ref_utils.PiecewiseLinear([0, 5, 24], [0, 0.4, 0])
    os_ = our_utils.PiecewiseLinear([0, 5, 24], [0, 0.4, 0])
    for t in np.linspace(0, 24, 97):
        worst = max(worst, abs(float(rs(t)) - float(os_(t))))
    if hasattr(ref_utils, "Exp") and hasattr(our_utils, "Exp"):
        re_, oe_ = ref_utils.Exp(2, 0.3, 5), our_utils.Exp(2, 0.3, 5)
        for t in np.linspace(0, 12, 49):
            worst = max(worst, abs(float(re_(t)) - float(oe_(t))))
    for name_ds in ("CIFAR10", "CIFAR100", "EMNIST", "ImageNet", "PERSONA"):
        try:
            want = ref_utils.num_classes_of_dataset(name_ds)
        except Exception:               # noqa: BLE001 -- a name the reference does not know
            continue
        worst = max(worst, float(want != our_utils.num_classes_of_dataset(name_ds)))
    ds6 = Sizes([4] * 10)
    for lbs, nw in ((-1, 3), (2, 3), (5, 4)):
        worst = max(worst, abs(float(ref_utils.steps_per_epoch(lbs, ds6, nw)) - float(our_utils.steps_per_epoch(lbs, ds6, nw))))
    v = torch.randn(500, generator=torch.Generator().manual_seed(9))
    worst = max(worst, float((ref_utils._topk(v.clone(), 37) - our_utils._topk(v.clone(), 37)).abs().max()))
    m1, m2 = make_model(3), make_model(3)
    worst = max(worst, float((ref_utils.get_param_vec(m1) - our_utils.get_param_vec(m2)).abs().max()))
    newv = torch.randn(n_par, generator=torch.Generator().manual_seed(10))
    ref_utils.set_param_vec(m1, newv.clone())
    our_utils.set_param_vec(m2, newv.clone())
    worst = max(worst, max(float((a - b).abs().max()) for a, b in zip(m1.parameters(), m2.parameters())))
    out["utils/schedules_topk_flat_vectors"] = worst
    print(json.dumps(out))
    return 0


if __name__ == "__main__":
    sys.exit(main())
