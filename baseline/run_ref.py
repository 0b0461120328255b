"""Benchmark harness for the UNMODIFIED reference (pursueorigin/CommEfficient), installed
under ``baseline/_ref`` by ``pip install --target``.  It drives the reference
through its own public API -- ``utils.parse_args``, ``models.ResNet9``,
``fed_aggregator.FedModel`` / ``FedOptimizer``, ``cv_train.compute_loss_train`` -- with
its stock parameter-server + worker-process + NCCL ``reduce`` code path.  None of the
commefficient_b200 models, kernels or engine are imported here.

What this harness adds around the reference (no reference file is edited):
  * ``baseline/shims`` on ``sys.path``/``PYTHONPATH``: stand-ins for the packages the reference
    imports but that do not exist on this image (csvec, fixup, pytorch_transformers, orjson)
    and two API-removal fixes (torchvision.models.utils, np.float);
  * synthetic CIFAR-10-shaped round batches built directly as tensors (the reference's PIL
    dataloader is kept out of the timed region for both arms);
  * CUDA-event timing around K rounds of ``model(batch); opt.step()`` on the PS process.

Run as a subprocess by ``bench.py --impl reference``; prints one JSON line.
"""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REF_PKG = os.path.join(HERE, "_ref", "CommEfficient")
SHIMS = os.path.join(HERE, "shims")


def _prepare_paths():
    for p in (SHIMS, REF_PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    extra = os.pathsep.join([REF_PKG, SHIMS])
    os.environ["PYTHONPATH"] = extra + os.pathsep + os.environ.get("PYTHONPATH", "")


def make_round(num_clients_round, per_client, num_total_clients, seed, torch):
    g = torch.Generator().manual_seed(seed)
    protos = torch.randn((10, 3, 32, 32), generator=torch.Generator().manual_seed(1234))
    n = num_clients_round * per_client
    labels = torch.randint(0, 10, (n,), generator=g)
    images = protos[labels] + 0.5 * torch.randn((n, 3, 32, 32), generator=g)
    first = (seed * num_clients_round) % max(1, num_total_clients - num_clients_round)
    ids = (torch.arange(num_clients_round).repeat_interleave(per_client) + first).long()
    return ids, images, labels


def main():
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--clients_per_round", type=int, default=100)
    ap.add_argument("--per_client", type=int, default=5)
    ap.add_argument("--num_clients", type=int, default=10000)
    ap.add_argument("--port", type=int, default=5315)
    cfg = ap.parse_args()

    _prepare_paths()
    if cfg.gpus == 1:
        os.environ["CEB_REF_ONE_GPU"] = "1"
    import _ref_compat  # noqa: F401
    import torch
    import torch.multiprocessing as multiprocessing
    multiprocessing.set_start_method("spawn")

    ref_argv = ["--dataset_name", "CIFAR10", "--mode", "sketch", "--error_type", "virtual",
                "--local_momentum", "0", "--virtual_momentum", "0.9", "--num_workers", str(cfg.clients_per_round),
                "--num_clients", str(cfg.num_clients), "--local_batch_size", "-1", "--num_epochs", "1",
                "--num_devices", str(cfg.gpus), "--device", "cuda", "--port", str(cfg.port),
                "--lr_scale", "0.4"]
    if cfg.gpus == 1:
        ref_argv.append("--share_ps_gpu")
    sys.argv = ["cv_train.py"] + ref_argv

    import utils as ref_utils                 # reference modules (bare names, as the reference does)
    import models as ref_models
    import cv_train as ref_cv
    from fed_aggregator import FedModel, FedOptimizer

    args = ref_utils.parse_args()
    torch.manual_seed(args.seed)
    model = ref_models.ResNet9(do_batchnorm=False, num_classes=10, new_num_classes=None,
                               bn_bias_freeze=False, bn_weight_freeze=False,
                               channels={"prep": 64, "layer1": 128, "layer2": 256, "layer3": 512})
    opt = torch.optim.SGD(model.parameters(), lr=1)
    fed_model = FedModel(model, ref_cv.compute_loss_train, args, ref_cv.compute_loss_val)
    fed_opt = FedOptimizer(opt, args)
    for group in fed_opt.param_groups:
        group["lr"] = 0.05
    n_procs = len(fed_model.update_forward_grad_ps)
    processed = (cfg.clients_per_round // n_procs) * n_procs      # reference drops the remainder (Q1)

    fed_model.train(True)
    rounds = [make_round(cfg.clients_per_round, cfg.per_client, cfg.num_clients, s, torch)
              for s in range(cfg.warmup + cfg.steps)]
    h2d = sum(t.numel() * t.element_size() for t in rounds[0][1:])
    for r in rounds[:cfg.warmup]:
        fed_model(r)
        fed_opt.step()
    torch.cuda.synchronize()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    start.record()
    loss = None
    for r in rounds[cfg.warmup:]:
        out = fed_model(r)
        fed_opt.step()
        loss = float(out[0].mean())
    stop.record()
    torch.cuda.synchronize()
    ms = start.elapsed_time(stop)
    wall_ms = (time.time() - t0) * 1e3
    ms = max(ms, wall_ms)          # PS stream is idle while workers compute: events == wall clock
    fed_model.finalize()
    ms_per_step = ms / cfg.steps
    print("REF_RESULT " + json.dumps({
        "ms_per_step": ms_per_step, "clients_processed_per_round": processed,
        "clients_per_sec": processed / (ms_per_step / 1e3), "worker_procs": n_procs,
        # per round the reference also copies the full weights host->device once per worker
        # process (fed_worker.py:41) and the dense update device->host (fed_aggregator.py:455)
        "h2d_bytes_per_step": h2d + 4 * model_numel(model) * n_procs,
        "d2h_bytes_per_step": 4 * model_numel(model),
        "last_loss": loss, "one_gpu_transport": os.environ.get("CEB_REF_ONE_GPU") == "1"}), flush=True)


def model_numel(model):
    return sum(p.numel() for p in model.parameters())


if __name__ == "__main__":
    main()
