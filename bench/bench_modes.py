#!/usr/bin/env python
"""Clients/sec of the other BASELINE.json configurations (ours): ResNet-18 / CIFAR-100-shaped
true_topk and local_topk with error accumulation; FEMNIST-shaped ResNet-9 (1x28x28, 62 classes,
3500 clients) fedavg vs uncompressed vs sketch.  Synthetic tensors, device-timed, max over ranks.

  python bench/bench_modes.py                       (1 GPU)
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/bench_modes.py
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from commefficient_b200 import models  # noqa: E402
from commefficient_b200.cv_train import compute_loss_train, compute_loss_val  # noqa: E402
from commefficient_b200.data_utils import synthetic_image_batch  # noqa: E402
from commefficient_b200.fed_aggregator import FedModel, FedOptimizer  # noqa: E402
from commefficient_b200.parallel.runtime import init_runtime  # noqa: E402
from commefficient_b200.utils import make_args  # noqa: E402

CONFIGS = [
    # name, model, model kwargs, image shape, classes, clients/gpu, images/client, args
    ("resnet18_c100_true_topk", "ResNet18", dict(num_classes=100), (3, 32, 32), 100, 16, 8,
     dict(mode="true_topk", error_type="virtual", local_momentum=0.0, virtual_momentum=0.9, k=50000)),
    ("resnet18_c100_local_topk_err", "ResNet18", dict(num_classes=100), (3, 32, 32), 100, 16, 8,
     dict(mode="local_topk", error_type="local", local_momentum=0.9, virtual_momentum=0.0, k=50000)),
    ("femnist_resnet9_sketch", "ResNet9", dict(num_classes=62, initial_channels=1), (1, 28, 28), 62, 32, 20,
     dict(mode="sketch", error_type="virtual", local_momentum=0.0, virtual_momentum=0.9)),
    ("femnist_resnet9_uncompressed", "ResNet9", dict(num_classes=62, initial_channels=1), (1, 28, 28), 62, 32, 20,
     dict(mode="uncompressed", local_momentum=0.0, virtual_momentum=0.9)),
    ("femnist_resnet9_fedavg", "ResNet9", dict(num_classes=62, initial_channels=1), (1, 28, 28), 62, 32, 20,
     dict(mode="fedavg", local_momentum=0.0, num_fedavg_epochs=2, fedavg_batch_size=10)),
]


def run_config(rt, name, model_name, mkw, shape, classes, cpg, ipc, akw, steps, warmup):
    world = rt.world_size
    W = cpg * world
    args = make_args(num_workers=W, num_clients=3500, local_batch_size=-1, dataset_name="EMNIST", device="cuda",
                     backend="auto", **akw)
    torch.manual_seed(0)
    model = getattr(models, model_name)(**mkw)
    fed = FedModel(model, compute_loss_train, args, compute_loss_val, runtime=rt)
    opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args)
    opt.param_groups[0]["lr"] = 0.05
    opt.step()                                   # publishes the LR (fedavg)
    fed.train(True)
    rounds = [synthetic_image_batch(W, ipc, num_classes=classes, shape=shape, seed=s, first_client=(s * W) % (3500 - W),
                                    pin=True) for s in range(3)]

    def one(i):
        out = fed(rounds[i % 3])
        opt.step()
        return out

    for i in range(warmup):
        one(i)
    rt.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(steps):
        out = one(warmup + i)
    b.record()
    rt.barrier()
    torch.cuda.synchronize()
    ms = rt.max_over_ranks(a.elapsed_time(b)) / steps
    res = {"config": name, "n_gpus": world, "clients_per_round": W, "images_per_client": ipc, "d": fed.fp.numel,
           "ms_per_round": ms, "clients_per_sec": W / (ms / 1e3), "client_batching": fed.worker.can_batch_clients(),
           "backend": rt.backend, "fused_reduce": fed.fused_reduce, "last_loss": float(out[0].mean())}
    for buf in (rt.peer.buffers if rt.peer is not None else []):
        pass
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--only", default="")
    cfg = ap.parse_args()
    rt = init_runtime(device="cuda", backend="auto")
    out = []
    for c in CONFIGS:
        if cfg.only and cfg.only not in c[0]:
            continue
        res = run_config(rt, *c, cfg.steps, cfg.warmup)
        if rt.rank == 0:
            print(json.dumps(res), flush=True)
        out.append(res)
    rt.shutdown()


if __name__ == "__main__":
    main()
