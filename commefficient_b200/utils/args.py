"""Command-line / config surface.

Keeps every flag name, ``dest`` and default of the reference parser
(/root/reference/CommEfficient/utils.py:102-230) so existing launch scripts
keep working, and adds B200-specific knobs with safe defaults.  The flag set is
table driven (``_FLAGS``) so the same table feeds ``parse_args`` (CLI) and
``make_args`` (programmatic construction used by tests and ``bench.py``).

Differences from the reference, all deliberate:
  * cross-flag validation raises ``ValueError`` with a message instead of a bare
    ``assert`` and also rejects the combinations that the reference only trips
    over deep inside a worker (sketch + local momentum, sketch + ``error_type
    none`` which yields an all-zero update forever -- Q8);
  * ``--num_devices`` means "ranks" (one per GPU): ``python cv_train.py --num_devices N`` starts N ranks itself
    unless torchrun already did (``parallel/runtime.py:run_driver``); there is no parameter-server
    GPU, so ``--share_ps_gpu`` is accepted and ignored;
  * ``--port`` probing is done by the launcher, not at parse time.
"""
from __future__ import annotations

import argparse
from types import SimpleNamespace
from typing import Any, Dict, Iterable, Optional

MODES = ("sketch", "true_topk", "local_topk", "fedavg", "uncompressed")
ERROR_TYPES = ("none", "local", "virtual")
DP_MODES = ("worker", "server")
DATASET_CLASSES = {"CIFAR10": 10, "CIFAR100": 100, "EMNIST": 62,
                   "ImageNet": 1000, "PERSONA": -1}


def num_classes_of_dataset(name: str) -> int:
    return DATASET_CLASSES[name]


def _flag(name, default=None, type=None, dest=None, choices=None, store_true=False, help=""):
    return dict(name=name, default=default, type=type, dest=dest, choices=choices,
                store_true=store_true, help=help)


# (group, flags) -- groups mirror the reference parser's comment blocks.
_FLAGS = [
    ("meta", [
        _flag("test", store_true=True, dest="do_test", help="one-round smoke mode"),
        _flag("mode", "sketch", str, choices=MODES),
        _flag("tensorboard", store_true=True, dest="use_tensorboard"),
        _flag("seed", 21, int),
    ]),
    ("data/model", [
        _flag("model", "ResNet9", str),
        _flag("finetune", store_true=True, dest="do_finetune"),
        _flag("checkpoint", store_true=True, dest="do_checkpoint"),
        _flag("checkpoint_path", "./checkpoint", str),
        _flag("finetune_path", "./finetune", str),
        _flag("finetuned_from", None, str, choices=tuple(DATASET_CLASSES)),
        _flag("num_results_train", 2, int),
        _flag("num_results_val", 2, int),
        _flag("dataset_name", "", str, choices=("",) + tuple(DATASET_CLASSES)),
        _flag("dataset_dir", "./dataset", str),
        _flag("batchnorm", store_true=True, dest="do_batchnorm"),
        _flag("nan_threshold", 999.0, float),
    ]),
    ("compression", [
        _flag("k", 50000, int),
        _flag("num_cols", 500000, int),
        _flag("num_rows", 5, int),
        _flag("num_blocks", 20, int),
        _flag("topk_down", store_true=True, dest="do_topk_down"),
    ]),
    ("optimisation", [
        _flag("local_momentum", 0.9, float),
        _flag("virtual_momentum", 0.0, float),
        _flag("weight_decay", 5e-4, float),
        _flag("num_epochs", 24.0, float),
        _flag("num_fedavg_epochs", 1, int),
        _flag("fedavg_batch_size", -1, int),
        _flag("fedavg_lr_decay", 1.0, float),
        _flag("error_type", "none", str, choices=ERROR_TYPES),
        _flag("lr_scale", None, float),
        _flag("pivot_epoch", 5.0, float),
    ]),
    ("parallelisation", [
        _flag("port", 5315, int),
        _flag("num_clients", None, int),
        _flag("num_workers", 1, int, help="clients sampled per round"),
        _flag("device", None, str, choices=("cpu", "cuda")),
        _flag("num_devices", 1, int, help="ranks (one per GPU)"),
        _flag("share_ps_gpu", store_true=True, help="accepted, ignored (no PS GPU)"),
        _flag("iid", store_true=True, dest="do_iid"),
        _flag("train_dataloader_workers", 0, int),
        _flag("val_dataloader_workers", 0, int),
    ]),
    ("gpt2", [
        _flag("model_checkpoint", "gpt2", str),
        _flag("num_candidates", 2, int),
        _flag("max_history", 2, int),
        _flag("local_batch_size", 8, int),
        _flag("valid_batch_size", 8, int),
        _flag("microbatch_size", -1, int),
        _flag("lm_coef", 1.0, float),
        _flag("mc_coef", 1.0, float),
        _flag("max_grad_norm", None, float),
        _flag("personality_permutations", 1, int),
        _flag("eval_before_start", store_true=True),
    ]),
    ("dp", [
        _flag("dp", store_true=True, dest="do_dp"),
        _flag("dp_mode", "worker", str, choices=DP_MODES),
        _flag("l2_norm_clip", 1.0, float),
        _flag("noise_multiplier", 0.0, float),
    ]),
    # ---- additions of this framework (not in the reference) ----
    ("b200", [
        _flag("backend", "auto", str, choices=("auto", "b200", "nccl", "gloo"),
              help="collective backend: b200 = peer-memory kernels over NVLink"),
        _flag("precision", "tf32", str, choices=("fp32", "tf32", "bf16"),
              help="tf32 (default): fp32 activations, TF32 tensor-core inputs, as the reference under cuDNN; "
                   "bf16: fast path; fp32: IEEE fp32 on the library, no tensor cores"),
        _flag("replicated_tail", store_true=True,
              help="multi-GPU sketch / true_topk: every rank recomputes the whole server tail (unsketch + top-k) "
                   "instead of the sharded tail with the distributed select"),
        _flag("p2", 0, int,
              help="FetchSGD second round: re-rank p2*k sketch candidates by exact value"),
        _flag("synthetic", store_true=True, help="generate data of the dataset's shape"),
        _flag("compat_microbatch_sum", store_true=True,
              help="reproduce reference Q6: sum (not mean) over microbatches"),
        _flag("compat_nonpersistent_client_state", store_true=True,
              help="reproduce reference Q3: local momentum/error never persist on GPU"),
        _flag("no_client_batching", store_true=True,
              help="always run one backward per client (disables the linearity fast path)"),
        _flag("client_lanes", 8, int,
              help="per-client modes (local_topk, fedavg, clipping, DP, topk_down): number of clients simulated "
                   "concurrently, each on its own model replica and CUDA stream (1 = one after the other)"),
        _flag("replica_check_every", 0, int,
              help="multi-GPU: every N rounds compare a checksum of the weight replicas across ranks and abort on a "
                   "mismatch (0 = never); the SPMD design never re-broadcasts weights, so drift would be silent"),
        _flag("resume", None, str, help="path of a *.fedstate.pt sidecar to resume from"),
        _flag("watchdog_timeout", 600.0, float, help="seconds before a hung round aborts all ranks"),
        _flag("flat_layout", "auto", str, choices=("auto", "reference", "nhwc"),
              help="coordinate order of conv weights in the flat buffer: reference = (out,in,kh,kw); "
                   "nhwc = (out,kh,kw,in), channels_last end to end (auto: nhwc on CUDA)"),
        _flag("no_cuda_graphs", store_true=True,
              help="do not capture the batched client forward/backward into a CUDA graph"),
        _flag("no_sketch_overlap", store_true=True,
              help="sketch the whole gradient after the backward pass instead of sketching finished ranges "
                   "on a second stream while the backward is still running"),
        _flag("defer_results", store_true=True,
              help="return each round's losses / byte counts as lazily materialised arrays (async D2H): "
                   "call opt.step() before reading them and the host never stalls the GPU"),
        _flag("trace_phases", "", str, help="append per-phase CUDA-event timings (max over ranks) to this "
                                            "JSONL file; phases are NVTX ranges too"),
        _flag("trace_every", 50, int, help="rounds per --trace_phases record"),
        _flag("no_native_nn", store_true=True,
              help="keep conv/linear layers on cuDNN/cuBLAS instead of the hand-written tcgen05 kernels"),
    ]),
]


def iter_flags() -> Iterable[Dict[str, Any]]:
    for _group, flags in _FLAGS:
        for f in flags:
            yield f


def _dest(f) -> str:
    return f["dest"] or f["name"]


def build_parser(default_lr: Optional[float] = None, model_names: Optional[Iterable[str]] = None):
    parser = argparse.ArgumentParser(description="commefficient_b200")
    for group, flags in _FLAGS:
        g = parser.add_argument_group(group)
        for f in flags:
            kw = dict(dest=_dest(f), help=f["help"] or None)
            if f["store_true"]:
                kw["action"] = "store_true"
            else:
                kw["type"] = f["type"]
                kw["default"] = f["default"]
                if f["choices"] is not None:
                    kw["choices"] = f["choices"]
            if f["name"] == "lr_scale":
                kw["default"] = default_lr
            if f["name"] == "model" and model_names is not None:
                kw["choices"] = tuple(model_names)
            g.add_argument("--" + f["name"], **kw)
    return parser


def default_device() -> str:
    import torch
    return "cuda" if torch.cuda.is_available() else "cpu"


def validate_args(args) -> None:
    """Cross-flag rules (reference: utils.py:225-228, fed_worker.py:221-228,
    fed_aggregator.py:484-486,512,545,573-576) checked up front."""
    def bad(msg):
        raise ValueError("invalid flag combination: " + msg)

    if args.mode == "fedavg":
        if args.local_batch_size != -1:
            bad("fedavg needs --local_batch_size -1 (whole client dataset per round)")
        if args.local_momentum != 0:
            bad("fedavg needs --local_momentum 0")
        if args.error_type != "none":
            bad("fedavg needs --error_type none")
    if args.mode == "sketch":
        if args.local_momentum != 0:
            bad("sketch mode cannot apply momentum factor masking per client; "
                "use --local_momentum 0 (momentum lives in the server sketch)")
        if args.error_type == "local":
            bad("sketch mode cannot keep a per-client error (the sketch does not say "
                "which part of it is error); use --error_type virtual")
        if args.error_type == "none":
            bad("sketch mode with --error_type none never writes the error sketch, so "
                "every update is exactly zero (reference quirk Q8); use --error_type virtual")
    if args.mode == "uncompressed" and args.error_type == "local":
        bad("uncompressed mode has nothing to feed back; use --error_type none")
    if args.mode == "true_topk" and args.error_type != "virtual":
        bad("true_topk needs --error_type virtual")
    if args.mode == "local_topk" and args.error_type not in ("local", "none"):
        bad("local_topk needs --error_type local or none")
    if args.mode in ("sketch", "true_topk", "local_topk") and args.k <= 0:
        bad("--k must be positive")
    if args.num_workers <= 0:
        bad("--num_workers must be positive")


def finalize_args(args):
    if getattr(args, "device", None) is None:
        args.device = default_device()
    validate_args(args)
    return args


def parse_args(default_lr: Optional[float] = None, argv=None, model_names=None):
    """Parse the CLI.  ``model_names`` defaults to every capitalised export of
    ``commefficient_b200.models`` (reference: utils.py:114-118)."""
    if model_names is None:
        from .. import models as _models
        model_names = _models.model_names()
    args = build_parser(default_lr, model_names).parse_args(argv)
    return finalize_args(args)


def make_args(**overrides):
    """Programmatic equivalent of ``parse_args`` (defaults + overrides by dest)."""
    ns = SimpleNamespace()
    for f in iter_flags():
        setattr(ns, _dest(f), False if f["store_true"] else f["default"])
    known = set(vars(ns))
    for key, val in overrides.items():
        if key not in known:
            raise KeyError("unknown argument %r" % key)
        setattr(ns, key, val)
    return finalize_args(ns)
