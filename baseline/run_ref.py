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


def make_round(num_clients_round, per_client, num_total_clients, seed, torch, noise=0.5):
    g = torch.Generator().manual_seed(seed)
    protos = torch.randn((10, 3, 32, 32), generator=torch.Generator().manual_seed(1234))
    n = num_clients_round * per_client
    labels = torch.randint(0, 10, (n,), generator=g)
    images = protos[labels] + noise * torch.randn((n, 3, 32, 32), generator=g)
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
    ap.add_argument("--config", default="resnet9_sketch", help="a key of baseline/bench_configs.py:CONFIGS")
    # accuracy-vs-round curve mode (bench/curves.py): same data, initial weights and LR schedule as our arm
    ap.add_argument("--curve_out", default="")
    ap.add_argument("--curve_init", default="", help="state_dict file with the initial weights")
    ap.add_argument("--noise", type=float, default=0.5)
    ap.add_argument("--lr_peak", type=float, default=0.1)
    ap.add_argument("--pivot", type=int, default=0)
    ap.add_argument("--val_every", type=int, default=25)
    ap.add_argument("--val_images", type=int, default=1000)
    ap.add_argument("--k", type=int, default=50000)
    ap.add_argument("--num_cols", type=int, default=500000)
    cfg = ap.parse_args()

    _prepare_paths()
    sys.path.insert(0, HERE)
    import bench_configs
    if bench_configs.CONFIGS[cfg.config]["family"] == "gpt2":
        os.environ["CEB_REF_GPT2"] = "1"          # the pytorch_transformers stand-in builds its GPT-2 adapter classes
    if cfg.gpus == 1:
        os.environ["CEB_REF_ONE_GPU"] = "1"
    import _ref_compat  # noqa: F401
    import torch
    import torch.multiprocessing as multiprocessing
    multiprocessing.set_start_method("spawn")

    bc = bench_configs.CONFIGS[cfg.config]
    gpt2 = bc["family"] == "gpt2"
    if bc["family"] == "cv" and bc["model"] == "ResNet9" and tuple(bc["shape"])[1] != 32:
        # the reference's ResNet-9 ends in MaxPool2d(4) (models/resnet9.py:95): on 28 x 28 input the map is 3 x 3 by
        # then and the layer raises ("Output size is too small") -- its worker dies and the server waits 600 s
        print("REF_UNAVAILABLE the reference's ResNet9 raises on %dx%d input (MaxPool2d(4) over a 3x3 map); see the "
              "femnist32_* configurations for the same task on 32x32-padded images" % tuple(bc["shape"])[1:], flush=True)
        return
    flags = dict(bc["flags"])
    ref_argv = ["--dataset_name", bc["dataset"], "--mode", flags["mode"], "--error_type", flags.get("error_type", "none"),
                "--local_momentum", str(flags.get("local_momentum", 0.0)),
                "--virtual_momentum", str(flags.get("virtual_momentum", 0.0)), "--num_workers", str(cfg.clients_per_round),
                "--num_clients", str(cfg.num_clients), "--local_batch_size", str(bc["per_client"] if gpt2 else -1),
                "--num_epochs", "1", "--num_devices", str(cfg.gpus), "--device", "cuda", "--port", str(cfg.port),
                "--lr_scale", "0.4", "--k", str(flags.get("k", cfg.k)), "--num_cols", str(cfg.num_cols),
                "--valid_batch_size", "500", "--weight_decay", str(flags.get("weight_decay", 5e-4))]
    if flags["mode"] == "fedavg":
        ref_argv += ["--num_fedavg_epochs", str(flags["num_fedavg_epochs"]), "--fedavg_batch_size",
                     str(flags["fedavg_batch_size"])]
    if gpt2:
        ref_argv += ["--num_results_train", "1", "--num_candidates", str(bc["candidates"])]
    else:
        ref_argv += ["--model", bc["model"]]
    if cfg.gpus == 1:
        ref_argv.append("--share_ps_gpu")
    sys.argv = ["cv_train.py"] + ref_argv

    import utils as ref_utils                 # reference modules (bare names, as the reference does)
    import models as ref_models
    from fed_aggregator import FedModel, FedOptimizer

    args = ref_utils.parse_args(default_lr=4e-2) if gpt2 else ref_utils.parse_args()
    torch.manual_seed(args.seed)
    if gpt2:
        import gpt2_train as ref_train
        from pytorch_transformers import GPT2DoubleHeadsModel
        model = GPT2DoubleHeadsModel.from_pretrained("gpt2")          # random init (no network), HF transformers
        model.resize_token_embeddings(50257 + 5)
        args.len_tokenizer = 50257 + 5
        loss_train, loss_val = ref_train.compute_loss_train, ref_train.compute_loss_val
    else:
        import cv_train as ref_train
        if bc["model"] == "ResNet9":
            kw = dict(bc["model_kw"])
            model = ref_models.ResNet9(do_batchnorm=False, num_classes=kw.get("num_classes", 10), new_num_classes=None,
                                       bn_bias_freeze=False, bn_weight_freeze=False,
                                       initial_channels=kw.get("initial_channels", 3),
                                       channels={"prep": 64, "layer1": 128, "layer2": 256, "layer3": 512})
        else:
            model = getattr(ref_models, bc["model"])(**bc["model_kw"])
        loss_train, loss_val = ref_train.compute_loss_train, ref_train.compute_loss_val
    if cfg.curve_init:
        model.load_state_dict(torch.load(cfg.curve_init, map_location="cpu"))
    opt = torch.optim.SGD(model.parameters(), lr=1)
    fed_model = FedModel(model, loss_train, args, loss_val)
    fed_opt = FedOptimizer(opt, args)
    for group in fed_opt.param_groups:
        group["lr"] = 0.01 if gpt2 else 0.05
    if flags["mode"] == "fedavg":
        fed_opt.step()                        # publishes the learning rate to the workers (fed_aggregator.py:441-444)
    n_procs = len(fed_model.update_forward_grad_ps)
    processed = (cfg.clients_per_round // n_procs) * n_procs      # reference drops the remainder (Q1)

    if cfg.curve_out:
        return run_curve(cfg, fed_model, fed_opt, torch)

    fed_model.train(True)
    make = bench_configs.gpt2_round if gpt2 else bench_configs.cv_round
    rounds = [make(bc, cfg.clients_per_round, s, torch) for s in range(cfg.warmup + cfg.steps)]
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


def curve_lr(t, rounds, pivot, peak):
    """Triangular schedule of the reference's cv_train (0 -> peak at `pivot` -> 0 at `rounds`), at round t >= 1."""
    import numpy as np
    return float(np.interp(t, [0, pivot, rounds], [0.0, peak, 0.0]))


def run_curve(cfg, fed_model, fed_opt, torch):
    """Train for cfg.steps rounds on the shared synthetic task and record the per-round mean client loss /
    accuracy and the validation loss / accuracy every cfg.val_every rounds."""
    import numpy as np
    pivot = cfg.pivot or max(1, cfg.steps // 5)
    val = make_round(cfg.val_images // 5, 5, 10 ** 9, 999983, torch, cfg.noise)
    val = (torch.full_like(val[0], -1), val[1], val[2])
    rec = {"round": [], "train_loss": [], "train_acc": [], "val_round": [], "val_loss": [], "val_acc": []}
    for t in range(1, cfg.steps + 1):
        for group in fed_opt.param_groups:
            group["lr"] = curve_lr(t, cfg.steps, pivot, cfg.lr_peak)
        fed_model.train(True)
        out = fed_model(make_round(cfg.clients_per_round, cfg.per_client, cfg.num_clients, t, torch, cfg.noise))
        fed_opt.step()
        rec["round"].append(t)
        rec["train_loss"].append(float(np.mean(out[0])))
        rec["train_acc"].append(float(np.mean(out[1])))
        if t % cfg.val_every == 0 or t == cfg.steps:
            fed_model.train(False)
            vout = fed_model(val)
            rec["val_round"].append(t)
            rec["val_loss"].append(float(np.mean(vout[0])))
            rec["val_acc"].append(float(np.mean(vout[1])))
    fed_model.finalize()
    with open(cfg.curve_out, "w") as f:
        json.dump(rec, f)
    print("REF_CURVE " + cfg.curve_out, flush=True)


def model_numel(model):
    return sum(p.numel() for p in model.parameters())


if __name__ == "__main__":
    main()
