#!/usr/bin/env python
"""Accuracy-vs-round curves: ours (tf32 = matched precision, bf16 = fast path) against the UNMODIFIED
reference (``baseline/run_ref.py --curve_out``), ResNet-9 FetchSGD sketch mode on the same learnable
synthetic CIFAR-10-shaped task -- same prototypes, same per-round samples and client ids, same initial
weights (ours are saved and loaded into the reference model), same triangular learning-rate schedule
(/root/reference/CommEfficient/cv_train.py:393-404), same sketch geometry, momentum and weight decay.

  python bench/curves.py --rounds 300 --out profiles/curve_resnet9_sketch.json

BASELINE.json makes "matching accuracy-vs-round curves on synthetic data" part of the target; the file this
writes is what ``tests/test_gpu_curves.py`` checks a short version of.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def curve_lr(t, rounds, pivot, peak):
    import numpy as np
    return float(np.interp(t, [0, pivot, rounds], [0.0, peak, 0.0]))


def run_ours(cfg, precision, init_path=None, device="cuda"):
    import numpy as np
    import torch
    from commefficient_b200 import models
    from commefficient_b200.cv_train import compute_loss_train, compute_loss_val
    from commefficient_b200.data_utils import synthetic_image_batch
    from commefficient_b200.fed_aggregator import FedModel, FedOptimizer
    from commefficient_b200.utils import make_args

    args = make_args(mode="sketch", error_type="virtual", local_momentum=0.0, virtual_momentum=0.9,
                     num_workers=cfg.clients_per_round, num_clients=cfg.num_clients, local_batch_size=-1,
                     dataset_name="CIFAR10", device=device, precision=precision, k=cfg.k, num_cols=cfg.num_cols,
                     valid_batch_size=500)
    torch.manual_seed(args.seed)
    model = models.ResNet9(num_classes=10)
    if init_path is not None:
        if os.path.exists(init_path):
            model.load_state_dict(torch.load(init_path, map_location="cpu"))
        else:
            torch.save(model.state_dict(), init_path)
    fed = FedModel(model, compute_loss_train, args, compute_loss_val)
    opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args)
    pivot = cfg.pivot or max(1, cfg.rounds // 5)
    val = synthetic_image_batch(cfg.val_images // 5, 5, seed=999983, noise=cfg.noise)
    val = (torch.full_like(val[0], -1), val[1], val[2])
    rec = {"round": [], "train_loss": [], "train_acc": [], "val_round": [], "val_loss": [], "val_acc": []}
    for t in range(1, cfg.rounds + 1):
        opt.param_groups[0]["lr"] = curve_lr(t, cfg.rounds, pivot, cfg.lr_peak)
        fed.train(True)
        first = (t * cfg.clients_per_round) % max(1, cfg.num_clients - cfg.clients_per_round)
        out = fed(synthetic_image_batch(cfg.clients_per_round, cfg.per_client, seed=t, first_client=first,
                                        noise=cfg.noise))
        opt.step()
        rec["round"].append(t)
        rec["train_loss"].append(float(np.mean(out[0])))
        rec["train_acc"].append(float(np.mean(out[1])))
        if t % cfg.val_every == 0 or t == cfg.rounds:
            fed.train(False)
            vout = fed(val)
            rec["val_round"].append(t)
            rec["val_loss"].append(float(np.mean(vout[0])))
            rec["val_acc"].append(float(np.mean(vout[1])))
    fed.rt.shutdown()
    return rec


def run_reference(cfg, init_path):
    ref_dir = os.path.join(ROOT, "baseline", "_ref", "CommEfficient")
    if not os.path.exists(os.path.join(ref_dir, "fed_aggregator.py")):
        return None
    out = tempfile.mktemp(suffix=".json")
    env = {k: v for k, v in os.environ.items()
           if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT")}
    cmd = [sys.executable, os.path.join(ROOT, "baseline", "run_ref.py"), "--gpus", "1", "--steps", str(cfg.rounds),
           "--clients_per_round", str(cfg.clients_per_round), "--per_client", str(cfg.per_client),
           "--num_clients", str(cfg.num_clients), "--port", str(5400 + os.getpid() % 400),
           "--curve_out", out, "--curve_init", init_path, "--noise", str(cfg.noise), "--lr_peak", str(cfg.lr_peak),
           "--pivot", str(cfg.pivot or max(1, cfg.rounds // 5)), "--val_every", str(cfg.val_every),
           "--val_images", str(cfg.val_images), "--k", str(cfg.k), "--num_cols", str(cfg.num_cols)]
    res = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=3000)
    if res.returncode != 0 or not os.path.exists(out):
        sys.stderr.write("reference curve run failed:\n" + (res.stderr or res.stdout)[-2000:] + "\n")
        return None
    with open(out) as f:
        rec = json.load(f)
    os.remove(out)
    return rec


def compare(a, b):
    """Largest gaps between two curves (train loss smoothed over 10 rounds; validation pointwise)."""
    import numpy as np

    def smooth(x, w=10):
        x = np.asarray(x, dtype=np.float64)
        if len(x) < w:
            return x
        return np.convolve(x, np.ones(w) / w, mode="valid")
    la, lb = smooth(a["train_loss"]), smooth(b["train_loss"])
    return {"max_abs_train_loss_gap_smoothed": float(np.max(np.abs(la - lb))),
            "max_abs_val_loss_gap": float(np.max(np.abs(np.asarray(a["val_loss"]) - np.asarray(b["val_loss"])))),
            "max_abs_val_acc_gap": float(np.max(np.abs(np.asarray(a["val_acc"]) - np.asarray(b["val_acc"])))),
            "final_val_acc": [a["val_acc"][-1], b["val_acc"][-1]]}


def parse(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--rounds", type=int, default=300)
    ap.add_argument("--clients_per_round", type=int, default=16)
    ap.add_argument("--per_client", type=int, default=5)
    ap.add_argument("--num_clients", type=int, default=10000)
    ap.add_argument("--noise", type=float, default=0.5, help="per-pixel noise of the prototype + noise images")
    ap.add_argument("--lr_peak", type=float, default=0.15)
    ap.add_argument("--pivot", type=int, default=0, help="round of the LR peak (default rounds / 5)")
    ap.add_argument("--val_every", type=int, default=25)
    ap.add_argument("--val_images", type=int, default=1000)
    ap.add_argument("--k", type=int, default=50000)
    ap.add_argument("--num_cols", type=int, default=500000)
    ap.add_argument("--arms", default="tf32,bf16,reference")
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "curve_resnet9_sketch.json"))
    return ap.parse_args(argv)


def main(argv=None):
    cfg = parse(argv)
    arms = cfg.arms.split(",")
    init = tempfile.mktemp(suffix=".pt")
    result = {"config": {k: getattr(cfg, k) for k in ("rounds", "clients_per_round", "per_client", "noise", "lr_peak",
                                                      "val_every", "val_images", "k", "num_cols")},
              "task": "ResNet-9, FetchSGD sketch 5 x num_cols, virtual momentum 0.9 + virtual error, weight decay 5e-4, "
                      "triangular LR; images = class prototype + noise (synthetic_image_batch / run_ref.make_round)"}
    first = True
    for arm in arms:
        if arm == "reference":
            continue
        result["ours_" + arm] = run_ours(cfg, arm, init)
        first = False
    if "reference" in arms:
        if first:
            run_ours(argparse.Namespace(**{**vars(cfg), "rounds": 0}), "tf32", init)      # just writes the initial weights
        ref = run_reference(cfg, init)
        if ref is not None:
            result["reference"] = ref
        else:
            result["reference_unavailable"] = True
    if "reference" in result:
        for arm in arms:
            if "ours_" + arm in result:
                result["gap_%s_vs_reference" % arm] = compare(result["ours_" + arm], result["reference"])
    if "ours_tf32" in result and "ours_bf16" in result:
        result["gap_bf16_vs_tf32"] = compare(result["ours_bf16"], result["ours_tf32"])
    if os.path.exists(init):
        os.remove(init)
    os.makedirs(os.path.dirname(cfg.out), exist_ok=True)
    with open(cfg.out, "w") as f:
        json.dump(result, f)
    print(json.dumps({k: v for k, v in result.items() if k.startswith("gap_")}, indent=1))
    return result


if __name__ == "__main__":
    main()
