#!/usr/bin/env python
"""Where does a round's time go between kernels?  Times, with CUDA events, (a) the replay of the
batched-round CUDA graph alone, (b) the eager server step, (c) the whole round, on the headline
config (1 GPU).  Compare (a) with the per-kernel sum in profiles/launches_resnet9_sketch_1gpu_native.txt."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from commefficient_b200 import models  # noqa: E402
from commefficient_b200.cv_train import compute_loss_train, compute_loss_val  # noqa: E402
from commefficient_b200.data_utils import synthetic_image_batch  # noqa: E402
from commefficient_b200.fed_aggregator import FedModel, FedOptimizer  # noqa: E402
from commefficient_b200.utils import make_args  # noqa: E402


def main():
    args = make_args(mode="sketch", error_type="virtual", local_momentum=0.0, virtual_momentum=0.9, num_workers=100,
                     num_clients=10000, local_batch_size=-1, k=50000, num_cols=500000, num_rows=5, device="cuda",
                     dataset_name="CIFAR10", weight_decay=5e-4, defer_results=True)
    torch.manual_seed(0)
    model = models.ResNet9(num_classes=10)
    fed = FedModel(model, compute_loss_train, args, compute_loss_val)
    opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args)
    opt.param_groups[0]["lr"] = 0.05
    fed.train(True)
    batches = [tuple(t.cuda() if i else t for i, t in enumerate(synthetic_image_batch(100, 5, seed=s, first_client=100 * s)))
               for s in range(4)]
    w = fed.worker
    ev = {"g0": [], "g1": [], "s0": [], "s1": [], "r0": [], "r1": []}

    def mark(key):
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        ev[key].append(e)

    for i in range(40):
        if i >= 10:
            mark("r0")
        out = fed(batches[i % 4])
        if i >= 10:
            mark("s0")
        opt.step()
        if i >= 10:
            mark("s1")
        float(np.mean(out[0]))
        if i >= 10:
            mark("r1")
    # graph replay alone, back to back
    entry = next(e for e in w._graphs.values() if e.get("graph") is not None)
    torch.cuda.synchronize()
    for _ in range(20):
        mark("g0")
        entry["graph"].replay()
        mark("g1")
    torch.cuda.synchronize()
    med = lambda a, b: float(np.median([x.elapsed_time(y) for x, y in zip(ev[a], ev[b])]))
    print(json.dumps({"graph_replay_ms": med("g0", "g1"), "server_step_ms": med("s0", "s1"),
                      "round_call_to_step_ms": med("r0", "s0"), "round_total_ms": med("r0", "r1")}))
    fed.rt.shutdown()


if __name__ == "__main__":
    main()
