#!/usr/bin/env python
"""GPT-2 small (124 M parameters) on PersonaChat-shaped synthetic batches, FetchSGD sketch mode
with virtual momentum (BASELINE.json config 4), ours only: clients/sec per round on N GPUs.

  python bench/bench_gpt2.py [--clients_per_gpu 8] [--local_batch 8] [--seq 128]
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/bench_gpt2.py
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from commefficient_b200 import models  # noqa: E402
from commefficient_b200.fed_aggregator import FedModel, FedOptimizer  # noqa: E402
from commefficient_b200.gpt2_train import compute_loss_train, compute_loss_val  # noqa: E402
from commefficient_b200.ops import native  # noqa: E402
from commefficient_b200.parallel.runtime import init_runtime  # noqa: E402
from commefficient_b200.utils import make_args  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--clients_per_gpu", type=int, default=8)
    ap.add_argument("--local_batch", type=int, default=8)
    ap.add_argument("--seq", type=int, default=128)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--no_native_nn", action="store_true")
    ap.add_argument("--no_client_batching", action="store_true")
    ap.add_argument("--no_cuda_graphs", action="store_true")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "tf32", "fp32"])
    ap.add_argument("--no_sketch_overlap", action="store_true")
    ap.add_argument("--mode", default="sketch")
    cfg = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    W = cfg.clients_per_gpu * world
    args = make_args(mode=cfg.mode, error_type="virtual" if cfg.mode == "sketch" else "none", no_sketch_overlap=cfg.no_sketch_overlap, local_momentum=0.0, virtual_momentum=0.9, num_workers=W,
                     num_clients=17568, local_batch_size=cfg.local_batch, dataset_name="PERSONA", device="cuda",
                     num_results_train=1, no_native_nn=cfg.no_native_nn, no_client_batching=cfg.no_client_batching, no_cuda_graphs=cfg.no_cuda_graphs,
                     weight_decay=0.0, precision=cfg.precision)
    rt = init_runtime(args)
    torch.manual_seed(0)
    model = models.GPT2DoubleHeadsModel()
    model.resize_token_embeddings(50257 + 5)
    fed = FedModel(model, compute_loss_train, args, compute_loss_val)
    opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args)
    opt.param_groups[0]["lr"] = 0.01
    fed.train(True)
    g = torch.Generator().manual_seed(1)
    n = W * cfg.local_batch
    rounds = []
    for r in range(3):
        ids = torch.randint(0, 50257, (n, 2, cfg.seq), generator=g)
        labels = ids.clone()
        labels[:, 0] = -1
        labels[:, 1, : cfg.seq // 2] = -1
        cid = (torch.arange(W).repeat_interleave(cfg.local_batch) + r * W).long()
        rounds.append(tuple(t.pin_memory() if t.dtype != torch.long or t.dim() > 1 else t for t in
                            (cid, ids, torch.full((n, 2), cfg.seq - 1), labels, torch.randint(0, 2, (n,), generator=g),
                             torch.randint(0, 50257, (n, 2, cfg.seq), generator=g))))

    def one(i):
        out = fed(rounds[i % len(rounds)])
        opt.step()
        return out

    for i in range(cfg.warmup):
        one(i)
    rt.barrier()
    torch.cuda.synchronize()
    native.reset_launch_count()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(cfg.steps):
        out = one(cfg.warmup + i)
    b.record()
    rt.barrier()
    torch.cuda.synchronize()
    ms = rt.max_over_ranks(a.elapsed_time(b)) / cfg.steps
    if rt.rank == 0:
        print(json.dumps({"model": "GPT2DoubleHeads small", "d": fed.fp.numel, "mode": "sketch", "n_gpus": world,
                          "clients_per_round": W, "sequences_per_client": cfg.local_batch * 2, "seq_len": cfg.seq,
                          "ms_per_round": ms, "clients_per_sec": W / (ms / 1e3),
                          "tokens_per_sec": n * 2 * cfg.seq / (ms / 1e3), "native_nn": not cfg.no_native_nn, "dtype": cfg.precision, "cuda_graphs": not cfg.no_cuda_graphs,
                          "client_batching": fed.worker.can_batch_clients(), "native_launches": native.launch_count(),
                          "last_loss": float(out[0].mean())}))
    rt.shutdown()


if __name__ == "__main__":
    main()
