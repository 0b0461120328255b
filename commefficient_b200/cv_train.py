"""Computer-vision driver: ResNets on CIFAR-10/100, FEMNIST, ImageNet shaped data.

Same flags, epoch/round loop, skip rules, NaN abort, logging columns, LR schedule and
checkpoint format as /root/reference/CommEfficient/cv_train.py (train :85-168,
run_batches :171-252, get_data_loaders :254-287, __main__ :289-421).  Every rank runs
this script (SPMD); rank 0 prints.  Differences: ``--eval_before_start`` works (Q10),
fine-tuning does not dereference an undefined attribute, ``--synthetic`` writes data of
the dataset's shape when the real files are absent (no network here), and
``--lr_scale`` defaults to 0.4 (cifar10-fast) instead of crashing when omitted.
"""
from __future__ import annotations

import math
import os

import numpy as np
import torch
import torch.nn.functional as F
from torch.optim.lr_scheduler import LambdaLR
from torch.utils.data import DataLoader

from . import data_utils, models
from .fed_aggregator import FedModel, FedOptimizer
from .parallel.runtime import init_runtime
from .utils import (PiecewiseLinear, TableLogger, Timer, make_logdir, num_classes_of_dataset, parse_args,
                    steps_per_epoch, union)


# ---------------------------------------------------------------- loss callbacks
def compute_loss_ce(model, batch, args):
    """``(mean cross-entropy, accuracy)`` of one client batch (cv_train.py:67-72)."""
    images, targets = batch
    logits = model(images).float()
    return F.cross_entropy(logits, targets), (logits.argmax(dim=1) == targets).float().mean()


def _per_sample_ce(model, batch, args):
    """Per-sample ``(loss, correct)`` vectors -- lets the worker batch clients together."""
    images, targets = batch
    if hasattr(model, "ce_per_sample"):
        return model.ce_per_sample(images, targets)
    logits = model(images).float()
    return F.cross_entropy(logits, targets, reduction="none"), (logits.argmax(dim=1) == targets).float()


compute_loss_ce.per_sample = _per_sample_ce


def compute_loss_train(model, batch, args):
    return compute_loss_ce(model, batch, args)


compute_loss_train.per_sample = _per_sample_ce


def compute_loss_val(model, batch, args):
    return compute_loss_ce(model, batch, args)


# ---------------------------------------------------------------- loops
def run_batches(model, opt, lr_scheduler, loader, training, epoch_fraction, args, is_main=True):
    if not training and epoch_fraction != 1:
        raise ValueError("validation always covers the full set")
    if not 0 < epoch_fraction <= 1:
        raise ValueError("epoch_fraction must be in (0, 1], got %r" % (epoch_fraction,))
    model.train(training)
    losses, accs = [], []
    if not training:
        for batch in loader:
            if model.val_batch_size(batch) < args.valid_batch_size:
                model.skip_val_batch()
                continue
            loss, acc = model(batch)
            losses.extend(loss)
            accs.extend(acc)
            if args.do_test:
                break
        return np.mean(losses), np.mean(accs), None, None

    n_clients = loader.dataset.num_clients
    download = torch.zeros(n_clients)
    upload = torch.zeros(n_clients)
    spe = steps_per_epoch(args.local_batch_size, loader.dataset, args.num_workers)
    timer = Timer()
    for i, batch in enumerate(loader):
        if i > spe * epoch_fraction:
            break
        lr_scheduler.step()
        if lr_scheduler.get_last_lr()[0] == 0:
            opt.step()          # publishes the LR for fedavg (no-op otherwise)
        # (the round as a whole: with rank-sharded loading the batch is only this rank's share of it)
        round_clients, round_samples = model.round_shape(batch)
        if (round_clients < args.num_workers if args.local_batch_size == -1
                else round_samples < args.num_workers * args.local_batch_size):
            model.skip_round()  # not enough clients left in this epoch / incomplete round
            continue
        loss, acc, dl, ul = model(batch)
        deferred = getattr(args, "defer_results", False)
        if deferred:
            opt.step()          # queued behind the round on the GPU before the host looks at any value
        if np.any(np.isnan(loss)):
            if is_main:
                print("LOSS OF %s IS NAN, TERMINATING TRAINING" % np.mean(loss))
            return np.nan, np.nan, np.nan, np.nan
        download += dl
        upload += ul
        if not deferred:
            opt.step()
        losses.extend(loss)
        accs.extend(acc)
        if args.dataset_name == "EMNIST" and is_main:
            print("LR: {:0.5f}, Loss: {:0.5f}, Acc: {:0.5f}, Time: {:0.2f}".format(
                lr_scheduler.get_last_lr()[0], float(np.mean(loss)), float(np.mean(acc)), timer()))
        if args.do_test:
            break
    return np.mean(losses), np.mean(accs), download, upload


def train(model, opt, lr_scheduler, train_loader, test_loader, args, writer, loggers=(), timer=None,
          is_main=True):
    timer = timer or Timer()
    total_down = total_up = 0.0
    summary = None
    if args.eval_before_start:
        _, acc0, _, _ = run_batches(model, None, None, test_loader, False, 1, args, is_main)
        timer()
        if is_main:
            print("Test acc at epoch 0: {:0.4f}".format(acc0))
    n_epochs = math.ceil(args.num_epochs)
    for epoch in range(n_epochs):
        frac = args.num_epochs - epoch if epoch == n_epochs - 1 else 1
        train_loss, train_acc, down, up = run_batches(model, opt, lr_scheduler, train_loader, True, frac,
                                                      args, is_main)
        if isinstance(train_loss, float) and math.isnan(train_loss):
            if is_main:
                print("TERMINATING TRAINING DUE TO NAN LOSS")
            return summary
        train_time = timer()
        down_mb = down.sum().item() / 2 ** 20
        up_mb = up.sum().item() / 2 ** 20
        total_down += down_mb
        total_up += up_mb
        test_loss, test_acc, _, _ = run_batches(model, None, None, test_loader, False, 1, args, is_main)
        test_time = timer()
        lr = lr_scheduler.get_last_lr()[0]
        summary = union({"epoch": epoch + 1, "lr": lr},
                        {"train_time": train_time, "train_loss": float(train_loss), "train_acc": float(train_acc),
                         "test_loss": float(test_loss), "test_acc": float(test_acc),
                         "down (MiB)": round(down_mb), "up (MiB)": round(up_mb),
                         "total_time": timer.total_time})
        if is_main:
            for logger in loggers:
                logger.append(summary)
            if writer is not None:
                for tag, val in (("Loss/train", train_loss), ("Loss/test", test_loss), ("Acc/train", train_acc),
                                 ("Acc/test", test_acc), ("Time/train", train_time), ("Time/test", test_time),
                                 ("Time/total", timer.total_time), ("Lr", lr)):
                    writer.add_scalar(tag, val, epoch)
    if is_main:
        nc = train_loader.dataset.num_clients
        print("Total Download (MiB): {:0.2f}".format(total_down))
        print("Total Upload (MiB): {:0.2f}".format(total_up))
        print("Avg Download Per Client: {:0.2f}".format(total_down / nc))
        print("Avg Upload Per Client: {:0.2f}".format(total_up / nc))
    return summary


# ---------------------------------------------------------------- data
def _ensure_synthetic(args):
    d = args.dataset_dir
    if os.path.exists(os.path.join(d, "stats.json")):
        return
    if args.dataset_name in ("CIFAR10", "CIFAR100"):
        ncls = num_classes_of_dataset(args.dataset_name)
        per_class = max(10, (args.num_clients or ncls * 10) * 5 // ncls)
        data_utils.write_synthetic_cifar(d, num_classes=ncls, images_per_class=per_class, num_val=1000,
                                         seed=args.seed)
    elif args.dataset_name == "EMNIST":
        data_utils.write_synthetic_emnist(d, num_clients=args.num_clients or 3500, seed=args.seed)
    elif args.dataset_name == "ImageNet":
        # one client per class (wnid), as in the real tree; a handful of small JPEGs per class
        data_utils.write_synthetic_imagenet(d, num_classes=args.num_clients or 8, images_per_class=4, val_per_class=2,
                                            seed=args.seed)
        # (FedImageNet derives stats.json from the tree on first use: do that once, here, before the other ranks look)
        data_utils.FedImageNet(d, "ImageNet", None, train=True, download=False)
    else:
        raise ValueError("no synthetic generator for dataset %r" % args.dataset_name)


def get_data_loaders(args, rank=0):
    name = args.dataset_name
    tf_train, tf_val = {
        "ImageNet": ("imagenet_train_transforms", "imagenet_val_transforms"),
        "CIFAR10": ("cifar10_train_transforms", "cifar10_test_transforms"),
        "CIFAR100": ("cifar100_train_transforms", "cifar100_test_transforms"),
        "EMNIST": ("femnist_train_transforms", "femnist_test_transforms"),
    }[name]
    if getattr(args, "synthetic", False):
        if rank == 0:
            _ensure_synthetic(args)
        rt = init_runtime(args)
        rt.barrier()
    cls = getattr(data_utils, "Fed" + name)
    train_ds = cls(args.dataset_dir, name, getattr(data_utils, tf_train), args.do_iid, args.num_clients,
                   train=True, download=not getattr(args, "synthetic", False) and name != "ImageNet")
    test_ds = cls(args.dataset_dir, name, getattr(data_utils, tf_val), train=False, download=False)
    # every rank builds the identical round schedule from the shared seed
    sampler = data_utils.FedSampler(train_ds, args.num_workers, args.local_batch_size, seed=args.seed)
    train_loader = DataLoader(train_ds, batch_sampler=sampler, num_workers=args.train_dataloader_workers)
    # sequential validation batches of valid_batch_size * num_workers samples; a rank loads only its own shards
    val_sampler = data_utils.ShardedEvalSampler(len(test_ds), args.valid_batch_size * args.num_workers,
                                                args.valid_batch_size)
    test_loader = DataLoader(test_ds, batch_sampler=val_sampler, num_workers=args.val_dataloader_workers)
    return train_loader, test_loader


# ---------------------------------------------------------------- model / optimiser construction
def build_model_and_groups(args):
    if args.do_test:
        channels = {"prep": 1, "layer1": 1, "layer2": 1, "layer3": 1}
        args.num_cols, args.num_rows, args.k = 10, 1, 10
    else:
        channels = {"prep": 64, "layer1": 128, "layer2": 256, "layer3": 512}
    if args.do_finetune:
        num_classes = num_classes_of_dataset(args.finetuned_from)
        new_classes = num_classes_of_dataset(args.dataset_name)
    else:
        num_classes, new_classes = num_classes_of_dataset(args.dataset_name), None
    config = {"num_classes": num_classes}
    if args.model == "ResNet9":
        config.update(channels=channels, new_num_classes=new_classes, do_batchnorm=args.do_batchnorm,
                      bn_bias_freeze=args.do_finetune, bn_weight_freeze=args.do_finetune)
    if args.dataset_name == "EMNIST" and args.model in ("ResNet9", "ResNet18", "FixupResNet18", "FixupResNet9"):
        config["initial_channels"] = 1
    if args.model == "FixupResNet9" and not args.do_test:
        config["channels"] = channels
    model = getattr(models, args.model)(**config)

    named = list(model.named_parameters())
    if args.model.startswith("Fixup"):
        bias = [p for n, p in named if "bias" in n]
        scale = [p for n, p in named if "scale" in n]
        other = [p for n, p in named if "bias" not in n and "scale" not in n]
        groups = [{"params": bias, "lr": 0.1}, {"params": scale, "lr": 0.1}, {"params": other, "lr": 1}]
    elif args.do_finetune:
        path = os.path.join(args.finetune_path, args.model + ".pt")
        legacy = args.finetune_path + args.model + ".pt"     # reference's separator-less name (Q17)
        model.load_state_dict(torch.load(path if os.path.exists(path) else legacy, map_location="cpu"))
        for p in model.parameters():
            p.requires_grad = False
        groups = list(model.finetune_parameters())
    else:
        groups = list(model.parameters())
    return model, groups


def main(argv=None):
    args = parse_args(argv=argv)
    if args.lr_scale is None:
        args.lr_scale = 0.4
    rt = init_runtime(args)
    is_main = rt.rank == 0
    if is_main:
        print(args)
    timer = Timer()
    np.random.seed(args.seed)
    torch.random.manual_seed(args.seed)
    torch.backends.cudnn.benchmark = True

    model, groups = build_model_and_groups(args)
    train_loader, test_loader = get_data_loaders(args, rt.rank)
    if args.num_clients is None:
        args.num_clients = train_loader.dataset.num_clients
    base_opt = torch.optim.SGD(groups, lr=1)
    fed_model = FedModel(model, compute_loss_train, args, compute_loss_val)
    # each rank's loader decodes / augments only the clients that rank simulates
    fed_model.attach_sampler(train_loader.batch_sampler)
    fed_model.attach_val_sampler(test_loader.batch_sampler)
    opt = FedOptimizer(base_opt, args)
    if args.resume:
        fed_model.load_fedstate(args.resume, opt)

    schedule = PiecewiseLinear([0, args.pivot_epoch, args.num_epochs], [0, args.lr_scale, 0])
    spe = steps_per_epoch(args.local_batch_size, train_loader.dataset, args.num_workers)
    lr_scheduler = LambdaLR(opt, lr_lambda=lambda step: schedule(step / spe))
    if args.resume:
        fed_model.restore_scheduler(lr_scheduler)

    writer = None
    if args.use_tensorboard and is_main:
        from torch.utils.tensorboard import SummaryWriter
        writer = SummaryWriter(log_dir=make_logdir(args))
    if is_main:
        print("Finished initializing in {:.2f} seconds".format(timer()))

    summary = train(fed_model, opt, lr_scheduler, train_loader, test_loader, args, writer,
                    loggers=(TableLogger(),), timer=timer, is_main=is_main)
    fed_model.finalize()
    if args.do_checkpoint:
        os.makedirs(args.checkpoint_path, exist_ok=True)
        if is_main:
            torch.save(fed_model.state_dict(), os.path.join(args.checkpoint_path, args.model + ".pt"))
        # every rank writes its own sidecar (server-state shard + the client rows it owns)
        fed_model.save_fedstate(os.path.join(args.checkpoint_path, args.model + ".fedstate.pt"), opt, lr_scheduler)
    rt.shutdown()
    return summary


if __name__ == "__main__":
    main()
