#!/usr/bin/env python
"""Headline benchmark: clients/sec per federated round (BASELINE.json), ResNet-9 on
CIFAR-10-shaped synthetic data, FetchSGD sketch mode (5 x 500 000 sketch, k = 50 000,
virtual momentum 0.9, virtual error), 100 clients per round per GPU, 5 images per client.

  python bench.py --gpus 1 --steps K --warmup W                       (ours, 1 GPU)
  python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...   (ours, N GPUs)
  python bench.py --impl reference --gpus N ...                        (unmodified reference)

Prints ONE JSON line on rank 0.  Timing: CUDA events on the compute stream around exactly K
rounds, barrier + synchronize on both sides, max over ranks.  ``value`` is timed with the
round's inputs already on the device; ``e2e`` is the same public API call
(``model(batch); opt.step()``) with every round's inputs copied from pinned host memory and the
per-client results read back to the host, as a user runs it.  Scaling is weak: each GPU
simulates 100 clients per round, so N GPUs run 100*N clients per round.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

sys.path.insert(0, os.path.join(ROOT, "baseline"))
import bench_configs  # noqa: E402  (plain data shared with the reference harness)

METRIC = "clients/sec per FL round (device-timed, max over ranks)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--config", default="resnet9_sketch", choices=sorted(bench_configs.CONFIGS),
                    help="which BASELINE.json configuration (default: the headline ResNet-9 FetchSGD sketch)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: clients_per_gpu x N clients per round; strong: a fixed round for every N")
    ap.add_argument("--precision", default="tf32", choices=["tf32", "bf16", "fp32"],
                    help="tf32 (headline): fp32 activations, tcgen05 kind::tf32 -- the precision the reference's fp32 "
                         "convolutions get from cuDNN; bf16: the fast path (reported as the extra key \"bf16\")")
    ap.add_argument("--no_extra", action="store_true", help="skip the extra bf16 measurement")
    ap.add_argument("--set", action="append", default=[], metavar="FLAG=VALUE",
                    help="override a training flag of the configuration (ours only; e.g. --set max_grad_norm=1.0): the "
                         "JSON line then carries the overrides in config.overrides and is NOT a BASELINE number")
    ap.add_argument("--ref_timeout", type=int, default=1500, help="seconds before the reference arm is abandoned")
    ap.add_argument("--backend", default="auto")
    ap.add_argument("--no_native_nn", action="store_true",
                    help="run conv/linear layers on cuDNN/cuBLAS instead of the hand-written tcgen05 kernels")
    ap.add_argument("--sync_results", action="store_true",
                    help="block on every round's results inside FedModel.__call__ (reference behaviour) instead "
                         "of --defer_results")
    ap.add_argument("--no_cuda_graphs", action="store_true",
                    help="launch the round's kernels one by one (for ncu launch lists; a graph replay hides them)")
    ap.add_argument("--client_lanes", type=int, default=int(os.environ.get("CEB_CLIENT_LANES", "8")),
                    help="per-client modes: clients simulated concurrently per GPU (model replicas on their own streams)")
    ap.add_argument("--clients_per_gpu", type=int, default=0, help="override the configuration's clients per GPU")
    ap.add_argument("--phases", action="store_true", help="also print a per-phase breakdown to stderr")
    ap.add_argument("--profile", action="store_true",
                    help="profiling aid: device-resident phase only, wrapped in cudaProfilerStart/Stop "
                         "(use with ncu --profile-from-start off); the printed number is NOT a bench value")
    return ap.parse_args()


# ---------------------------------------------------------------------------- clocks
class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.proc = None
        self.path = "/tmp/ceb_clocks_%d.csv" % os.getpid()
        self.gpu_index = gpu_index

    def start(self):
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu_index), "--query-gpu=" + self.QUERY,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        try:
            for line in open(self.path):
                parts = [p.strip() for p in line.split(",")]
                if len(parts) < 9:
                    continue
                try:
                    sm.append(float(parts[1]))
                    mx.append(float(parts[2]))
                except ValueError:
                    continue
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"),
                                     parts[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            os.remove(self.path)
        except Exception:
            pass
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ---------------------------------------------------------------------------- reference arm
def run_reference(cfg):
    ref_dir = os.path.join(ROOT, "baseline", "_ref", "CommEfficient")
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0            # the reference spawns its own worker processes from rank 0
    if not os.path.exists(os.path.join(ref_dir, "fed_aggregator.py")):
        print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref is not installed (see DESIGN.md)"}))
        return 0
    env = dict(os.environ)
    # The reference does its own rendezvous; scrub everything torchrun exported (in particular
    # TORCHELASTIC_USE_AGENT_STORE, which would make its rank 0 wait for an agent-hosted store).
    for k in list(env):
        if k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT", "LOCAL_WORLD_SIZE",
                 "GROUP_RANK", "ROLE_RANK", "ROLE_NAME", "GROUP_WORLD_SIZE", "ROLE_WORLD_SIZE",
                 "OMP_NUM_THREADS", "NCCL_ASYNC_ERROR_HANDLING") or k.startswith(("TORCHELASTIC", "TORCH_NCCL")):
            env.pop(k, None)
    bc = dict(bench_configs.CONFIGS[cfg.config])
    if cfg.clients_per_gpu:
        bc["clients_per_gpu"] = cfg.clients_per_gpu
    clients = bench_configs.clients_per_round(bc, cfg.gpus, cfg.scaling)
    cmd = [sys.executable, os.path.join(ROOT, "baseline", "run_ref.py"), "--gpus", str(cfg.gpus),
           "--steps", str(cfg.steps), "--warmup", str(cfg.warmup), "--clients_per_round", str(clients),
           "--per_client", str(bc["per_client"]), "--num_clients", str(bc["num_clients"]), "--config", cfg.config,
           "--port", str(5315 + os.getpid() % 500)]
    sampler = ClockSampler(0)
    sampler.start()
    try:
        res = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                             timeout=cfg.ref_timeout)
    except subprocess.TimeoutExpired:
        sampler.stop()
        print(json.dumps({"impl": "reference", "unavailable": "reference run exceeded %d s" % cfg.ref_timeout}))
        return 0
    clocks = sampler.stop()
    unavailable = [l for l in res.stdout.splitlines() if l.startswith("REF_UNAVAILABLE ")]
    if unavailable:
        print(json.dumps({"impl": "reference", "config": cfg.config, "unavailable": unavailable[-1][len("REF_UNAVAILABLE "):]}))
        return 0
    line = [l for l in res.stdout.splitlines() if l.startswith("REF_RESULT ")]
    if res.returncode != 0 or not line:
        tail = (res.stderr or res.stdout).strip().splitlines()[-3:]
        print(json.dumps({"impl": "reference", "unavailable": "reference run failed: " + " | ".join(tail)[:400]}))
        return 0
    r = json.loads(line[-1][len("REF_RESULT "):])
    value = r["clients_per_sec"]
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "clients/s", "n_gpus": cfg.gpus,
        "steps": cfg.steps, "warmup": cfg.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": cfg.scaling, "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
        "config": {"name": cfg.config, "model": bc["model"], "mode": bc["flags"]["mode"], "num_rows": 5,
                   "num_cols": 500000, "k": bc["flags"].get("k", 50000),
                   "global_batch": clients * bc["per_client"], "clients_per_round": clients,
                   "clients_processed_per_round": r["clients_processed_per_round"],
                   "images_per_client": bc["per_client"], "seq_len": bc.get("seq"),
                   "precision_note": "the reference's only mode: fp32 tensors, TF32 convolutions inside cuDNN "
                                     "(torch.backends.cudnn.allow_tf32 defaults to True), fp32 matmuls",
                   "parallelism": "ps+%dworkers" % r["worker_procs"],
                   "l2": "inputs + activations exceed L2; fresh inputs every round",
                   "one_gpu_transport_gloo": r["one_gpu_transport"]},
        "e2e": {"value": value, "unit": "clients/s", "h2d_bytes_per_step": r["h2d_bytes_per_step"],
                "d2h_bytes_per_step": r["d2h_bytes_per_step"]},
        "gpu_launches": 0, "clocks": clocks, "last_loss": r["last_loss"],
    }
    print(json.dumps(out))
    return 0


# ---------------------------------------------------------------------------- our arm
def run_ours(cfg):
    import numpy as np
    import torch
    from commefficient_b200 import models
    from commefficient_b200.cv_train import compute_loss_train, compute_loss_val
    from commefficient_b200.data_utils import synthetic_image_batch
    from commefficient_b200.fed_aggregator import FedModel, FedOptimizer
    from commefficient_b200.ops import native
    from commefficient_b200.ops import nn as nnops
    from commefficient_b200.parallel.runtime import init_runtime
    from commefficient_b200.utils import make_args

    if not torch.cuda.is_available():
        print(json.dumps({"metric": METRIC, "error": "no CUDA device"}))
        return 1
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != cfg.gpus:
        if world == 1 and cfg.gpus > 1:
            # convenience: relaunch ourselves under torchrun
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(cfg.gpus),
                   "--master-addr", "127.0.0.1", "--master-port", str(29400 + os.getpid() % 500)] + sys.argv
            return subprocess.call(cmd)
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (cfg.gpus, world))
    bc = dict(bench_configs.CONFIGS[cfg.config])
    if cfg.clients_per_gpu:
        bc["clients_per_gpu"] = cfg.clients_per_gpu
    clients = bench_configs.clients_per_round(bc, cfg.gpus, cfg.scaling)
    gpt2 = bc["family"] == "gpt2"
    n_distinct = 3 if gpt2 else 4

    def pin(t):
        return t.pin_memory() if torch.cuda.is_available() else t
    if gpt2:
        from commefficient_b200.gpt2_train import compute_loss_train as loss_train, compute_loss_val as loss_val
        host_rounds = [tuple(pin(t) for t in bench_configs.gpt2_round(bc, clients, s, torch)) for s in range(n_distinct)]
    else:
        loss_train, loss_val = compute_loss_train, compute_loss_val
        host_rounds = [tuple(pin(t) for t in bench_configs.cv_round(bc, clients, s, torch)) for s in range(n_distinct)]
    per_rank = clients // cfg.gpus if clients % cfg.gpus == 0 else -(-clients // cfg.gpus)
    h2d = sum(t[0:1].numel() * t.element_size() * per_rank * bc["per_client"] for t in host_rounds[0][1:])
    state = {}

    def measure(precision, steps, warmup, e2e=True, profile=False):
        flags = dict(bc["flags"])
        for item in cfg.set:
            k, _, v = item.partition("=")
            try:
                flags[k] = json.loads(v)
            except ValueError:
                flags[k] = v
        args = make_args(num_workers=clients, num_clients=bc["num_clients"],
                         local_batch_size=bc["per_client"] if gpt2 else -1, num_epochs=1.0, dataset_name=bc["dataset"],
                         device="cuda", precision=precision, backend=cfg.backend, model=bc["model"],
                         no_native_nn=cfg.no_native_nn, no_cuda_graphs=cfg.no_cuda_graphs, client_lanes=cfg.client_lanes,
                         defer_results=not cfg.sync_results, num_results_train=1 if gpt2 else 2, **flags)
        rt = init_runtime(args)
        torch.manual_seed(args.seed)
        torch.backends.cudnn.benchmark = True
        model = getattr(models, bc["model"])(**bc["model_kw"])
        if gpt2:
            model.resize_token_embeddings(50257 + 5)
        fed = FedModel(model, loss_train, args, loss_val)
        opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args)
        opt.param_groups[0]["lr"] = 0.01 if gpt2 else 0.05
        opt.step()                                 # publishes the learning rate (fedavg clients read it)
        fed.train(True)
        dev = rt.device
        dev_rounds = [(r[0],) + tuple(t.to(dev) for t in r[1:]) for r in host_rounds]
        d2h = clients * (args.num_results_train + 1) * 4

        def one_round(batch, read=True):
            out = fed(batch)
            opt.step()
            if read:
                # the step's result is read on the host (results are deferred: the async D2H issued by
                # fed() has been overtaken by the server step's launches, not skipped)
                float(np.mean(out[0]))
            return out

        def timed(rounds, steps, warmup, read_every_step):
            for i in range(warmup):
                one_round(rounds[i % len(rounds)])
            rt.barrier()
            torch.cuda.synchronize()
            native.reset_launch_count()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            out = None
            for i in range(steps):
                # device-timed loop: the loss is read once, after the last step (a driver that logs every K
                # rounds); end-to-end loop: the loss of EVERY step is read on the host
                out = one_round(rounds[(warmup + i) % len(rounds)], read=read_every_step or i == steps - 1)
            b.record()
            rt.barrier()
            torch.cuda.synchronize()
            ms = rt.max_over_ranks(a.elapsed_time(b))
            return ms / steps, native.launch_count(), out

        if profile:
            for i in range(warmup):
                one_round(dev_rounds[i % n_distinct])
            torch.cuda.synchronize()
            torch.cuda.profiler.start()
            for i in range(steps):
                one_round(dev_rounds[i % n_distinct])
            torch.cuda.synchronize()
            torch.cuda.profiler.stop()
            return None
        ms_dev, launches_dev, out = timed(dev_rounds, steps, warmup, read_every_step=cfg.sync_results)
        res = {"ms_dev": ms_dev, "launches": launches_dev, "args": args, "rt": rt, "d2h": d2h,
               "client_batching": fed.worker.can_batch_clients(), "native_nn": nnops.native_enabled(),
               "backend": rt.backend}
        if e2e:
            ms_e2e, launches, out = timed(host_rounds, steps, max(3, warmup), read_every_step=True)
            res.update(ms_e2e=ms_e2e, launches=launches)
        res["last_loss"] = float(out[0].mean())
        if cfg.phases:      # every rank runs the rounds (they contain collectives); rank 0 prints
            _phase_breakdown(fed, opt, dev_rounds, torch, rt.rank == 0)
        state["fed"] = fed
        return res

    rank0 = int(os.environ.get("RANK", "0")) == 0
    sampler = ClockSampler(int(os.environ.get("LOCAL_RANK", "0")))
    if cfg.profile:
        measure(cfg.precision, cfg.steps, cfg.warmup, profile=True)
        print(json.dumps({"profile_run": True, "steps": cfg.steps}))
        return 0
    if rank0:
        sampler.start()
    main = measure(cfg.precision, cfg.steps, cfg.warmup)
    clocks = sampler.stop() if rank0 else None
    extra = None
    if cfg.precision == "tf32" and not cfg.no_extra and cfg.config == "resnet9_sketch":
        # the fast path, same config and timing rules (device-timed + end to end), reported beside the headline
        extra = measure("bf16", cfg.steps, cfg.warmup)
    rt, args = main["rt"], main["args"]
    if rt.rank == 0:
        value = clients / (main["ms_dev"] / 1e3)
        e2e = clients / (main["ms_e2e"] / 1e3)
        line = {
            "metric": METRIC, "value": value, "unit": "clients/s", "n_gpus": cfg.gpus, "steps": cfg.steps,
            "warmup": cfg.warmup, "ms_per_step": main["ms_dev"], "higher_is_better": True, "scaling": cfg.scaling,
            "vs_baseline": None, "dtype": cfg.precision, "data": "synthetic",
            "config": {"name": cfg.config, "model": bc["model"], "mode": args.mode, "num_rows": args.num_rows,
                       "num_cols": args.num_cols,
                       "k": args.k, "global_batch": clients * bc["per_client"], "clients_per_round": clients,
                       "images_per_client": bc["per_client"], "seq_len": bc.get("seq"),
                       "parallelism": "dp%d" % cfg.gpus, "backend": main["backend"],
                       "client_batching": main["client_batching"], "native_nn": main["native_nn"],
                       "client_lanes": cfg.client_lanes, **({"overrides": list(cfg.set)} if cfg.set else {}),
                       "precision_note": "tf32 = fp32 activations / weights, tcgen05 kind::tf32 MMAs with fp32 "
                                         "accumulation, fp32 head: what the reference's fp32 model gets from cuDNN "
                                         "(allow_tf32 defaults to True)",
                       "l2": "weights+grads+sketch+activations (>300 MB/round) exceed the 126 MB L2; "
                             "%d distinct input batches rotate" % n_distinct},
            "e2e": {"value": e2e, "unit": "clients/s", "ms_per_step": main["ms_e2e"], "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": main["d2h"]},
            "gpu_launches": int(main["launches"]), "clocks": clocks, "last_loss": main["last_loss"],
        }
        if extra is not None:
            line["bf16"] = {"value": clients / (extra["ms_dev"] / 1e3), "ms_per_step": extra["ms_dev"],
                            "e2e_value": clients / (extra["ms_e2e"] / 1e3), "e2e_ms_per_step": extra["ms_e2e"],
                            "gpu_launches": int(extra["launches"]), "last_loss": extra["last_loss"]}
        print(json.dumps(line))
    rt.shutdown()
    return 0


def _phase_breakdown(fed, opt, rounds, torch, report=True):
    """Per-phase device times of one round (stderr; not part of the JSON contract)."""
    import numpy as np
    ev = lambda: torch.cuda.Event(enable_timing=True)
    acc = {}
    for i in range(5):
        e = [ev() for _ in range(3)]
        e[0].record()
        fed(rounds[i % len(rounds)])
        e[1].record()
        opt.step()
        e[2].record()
        torch.cuda.synchronize()
        acc.setdefault("clients+sketch+reduce", []).append(e[0].elapsed_time(e[1]))
        acc.setdefault("server_step", []).append(e[1].elapsed_time(e[2]))
    if report:
        print("phase ms: " + ", ".join("%s=%.3f" % (k, float(np.median(v))) for k, v in acc.items()), file=sys.stderr)


def main():
    if os.environ.get("CEB_HANG_DUMP"):          # debugging aid: Python stacks of every rank after N seconds
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["CEB_HANG_DUMP"]), repeat=False, exit=True)
    cfg = parse()
    if cfg.impl == "reference":
        return run_reference(cfg)
    return run_ours(cfg)


if __name__ == "__main__":
    sys.exit(main())
