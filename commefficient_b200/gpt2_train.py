"""GPT-2 double-heads on PersonaChat-shaped data.

Flags, loss definitions, per-round logging, linear LR decay and the train -> save ->
validate flow follow /root/reference/CommEfficient/gpt2_train.py (compute_loss_val :77-86,
compute_loss_train :88-99, add_special_tokens_ :101-112, train_gpt2 :115-147, test_gpt2
:149-167, run_batches :169-253, train :255-313, get_data_loaders :315-355).  The model and
tokenizer are this package's own (``models.gpt2``, ``data_utils.SimpleTokenizer``) because
``pytorch_transformers`` and the GPT-2 vocabulary files are not available offline;
``--model_checkpoint`` accepts a directory written by ``save_pretrained``, ``gpt2`` (GPT-2
small, random init) or ``gpt2-tiny`` (2 layers x 64, for tests).  Every rank runs this
script; rank 0 logs.
"""
from __future__ import annotations

import math
import os

import numpy as np
import torch
import torch.nn.functional as F
from torch.optim import SGD
from torch.optim.lr_scheduler import LambdaLR
from torch.utils.data import DataLoader

from . import data_utils
from .data_utils import (ATTR_TO_SPECIAL_TOKEN, FedPERSONA, FedSampler, SimpleTokenizer, load_tokenizer,
                         personachat_collate_fn)
from .fed_aggregator import FedModel, FedOptimizer
from .models.gpt2 import GPT2Config, GPT2DoubleHeadsModel, OpenAIGPTDoubleHeadsModel
from .parallel.runtime import init_runtime
from .utils import Logger, PiecewiseLinear, TableLogger, Timer, make_logdir, parse_args, steps_per_epoch, union

logger = Logger()

TINY = dict(vocab_size=512, n_positions=256, n_ctx=256, n_embd=64, n_layer=2, n_head=4)


# ---------------------------------------------------------------- losses / metrics
def inference(model, batch, args):
    model.eval()
    with torch.no_grad():
        input_ids, mc_token_ids, lm_labels, mc_labels, token_type_ids = batch
        lm_logits, mc_logits, *_ = model(input_ids, token_type_ids=token_type_ids, mc_token_ids=mc_token_ids)
        flat_logits = lm_logits[..., :-1, :].reshape(-1, lm_logits.size(-1))
        flat_labels = lm_labels[..., 1:].reshape(-1)
        return (flat_logits, mc_logits), (flat_labels, mc_labels)


def accuracy(y_pred, y):
    return (torch.argmax(y_pred, dim=1) == y).float().mean()


def compute_loss_val(model, batch, args):
    """``(LM negative log-likelihood over non-ignored tokens, multiple-choice accuracy)``.  Models that
    offer ``lm_rows`` score only the labeled positions (same value: the mean over non-ignored tokens)
    instead of materialising ``[N*C*T, V]`` logits (reference: gpt2_train.py:55-86)."""
    if hasattr(model, "lm_rows"):
        model.eval()
        with torch.no_grad():
            input_ids, mc_token_ids, lm_labels, mc_labels, token_type_ids = batch
            _, mc_logits, hidden = model(input_ids, token_type_ids=token_type_ids, mc_token_ids=mc_token_ids,
                                         need_logits=False, return_hidden=True)
            rows, mask = model.lm_rows(hidden, lm_labels)
            nll = rows.sum() / mask.sum().clamp_min(1)
            return nll, accuracy(mc_logits, mc_labels)
    (lm_logits, mc_logits), (lm_labels, mc_labels) = inference(model, batch, args)
    nll = F.cross_entropy(lm_logits.float(), lm_labels, ignore_index=-1)
    return nll, accuracy(mc_logits, mc_labels)


def compute_loss_train(model, batch, args):
    """``lm_coef * lm_loss + mc_coef * mc_loss``; no metrics (returns a 1-tuple)."""
    input_ids, mc_token_ids, lm_labels, mc_labels, token_type_ids = batch
    lm_loss, mc_loss, *_ = model(input_ids, token_type_ids=token_type_ids, mc_token_ids=mc_token_ids,
                                 mc_labels=mc_labels, lm_labels=lm_labels, need_logits=False)
    return lm_loss * args.lm_coef + mc_loss * args.mc_coef,


def _per_sample_train(model, batch, args, seg, num_segments):
    """Per-sample values ``v_s`` whose sum over a client's samples is ``n_i * loss_i`` with
    ``loss_i = lm_coef * (mean LM loss over the client's target tokens) + mc_coef * (mean MC loss)``
    -- exactly what ``compute_loss_train`` returns for that client alone, so the worker can run all
    of a rank's clients in one forward/backward."""
    input_ids, mc_token_ids, lm_labels, mc_labels, token_type_ids = batch
    _, mc_logits, hidden = model(input_ids, token_type_ids=token_type_ids, mc_token_ids=mc_token_ids,
                                 need_logits=False, return_hidden=True)
    B = input_ids.shape[0]
    rows, mask = model.lm_rows(hidden, lm_labels)      # [B*C, T]; LM head only where labeled
    tok_loss = rows.view(B, -1)
    tok_cnt = mask.view(B, -1).sum(1).float()
    n_i = torch.zeros(num_segments, device=seg.device).index_add_(0, seg, torch.ones_like(tok_cnt))
    ntok_i = torch.zeros(num_segments, device=seg.device).index_add_(0, seg, tok_cnt).clamp_min(1.0)
    lm_part = tok_loss.sum(1) * (n_i / ntok_i)[seg]
    mc_part = F.cross_entropy(mc_logits.float(), mc_labels, reduction="none")
    return lm_part * args.lm_coef + mc_part * args.mc_coef,


_per_sample_train.needs_segments = True
compute_loss_train.per_sample = _per_sample_train


def add_special_tokens_(model, tokenizer):
    """Add the five dialogue tokens and grow the embedding matrix accordingly."""
    base = len(tokenizer.encoder)
    added = tokenizer.add_special_tokens(ATTR_TO_SPECIAL_TOKEN)
    if added > 0:
        model.resize_token_embeddings(new_num_tokens=base + added)


# ---------------------------------------------------------------- loops
def run_batches(model, opt, lr_scheduler, loader, args, timer, training, epoch=None, epoch_fraction=None,
                logger=None, writer=None, is_main=True):
    if not training and epoch_fraction != 1:
        raise ValueError("validation always covers the full set")
    if not 0 < epoch_fraction <= 1:
        raise ValueError("epoch_fraction must be in (0, 1], got %r" % (epoch_fraction,))
    model.train(training)
    spe = steps_per_epoch(args.local_batch_size, loader.dataset, args.num_workers)
    if not training:
        nlls, accs = [], []
        for i, batch in enumerate(loader):
            if i > 5 and args.do_test and i < spe - 5:
                model.skip_val_batch()
                continue
            nll, acc = model(batch)
            nlls.append(np.mean(nll))
            accs.append(np.mean(acc))
        return np.mean(nlls), np.mean(accs), np.exp(np.mean(nlls))

    n_clients = loader.dataset.num_clients
    download, upload = torch.zeros(n_clients), torch.zeros(n_clients)
    offset = epoch * spe
    losses = []
    for i, batch in enumerate(loader):
        if i > 2 and args.do_test and i < spe - 10:
            model.skip_round()
            continue
        if i > spe * epoch_fraction:
            break
        lr_scheduler.step()
        if lr_scheduler.get_last_lr()[0] == 0:
            opt.step()
        # (the round as a whole: with rank-sharded loading the batch is only this rank's share of it)
        round_clients, round_samples = model.round_shape(batch)
        if (round_clients < args.num_workers if args.local_batch_size == -1
                else round_samples < args.num_workers * args.local_batch_size):
            model.skip_round()
            continue
        loss, dl, ul = model(batch)
        opt.step()              # before the byte counts / loss are read: with --defer_results those are lazy
        download += dl
        upload += ul
        loss = float(np.mean(loss))
        losses.append(loss)
        train_time = timer()
        lr = lr_scheduler.get_last_lr()[0]
        row = union({"batch_idx": i + 1 + offset, "lr": lr},
                    {"train_time": train_time, "train_loss": loss, "total_time": timer.total_time,
                     "down (MiB)": round(dl.sum().item() / 2 ** 20), "up (MiB)": round(ul.sum().item() / 2 ** 20)})
        if is_main:
            if writer is not None:
                writer.add_scalar("training/loss", loss, i + offset)
                writer.add_scalar("Lr", lr, i + offset)
                writer.add_scalar("Time/train", train_time, i + offset)
            if logger is not None:
                logger.append(row)
    return np.mean(losses), download, upload


def test_gpt2(model, val_loader, args, logger=None, timer=None, writer=None, is_main=True):
    nll, acc, ppl = run_batches(model, None, None, val_loader, args, timer, training=False, epoch_fraction=1,
                                is_main=is_main)
    stats = {"val_nll": float(nll), "val_acc": float(acc), "val_ppl": float(ppl), "val_time": timer(),
             "total_time": timer.total_time}
    if is_main:
        if writer is not None:
            writer.add_scalar("validation/nll", nll)
            writer.add_scalar("validation/acc", acc)
            writer.add_scalar("validation/ppl", ppl)
        print()
        (logger or TableLogger()).append(stats)
    return stats


def train_gpt2(model, opt, scheduler, train_loader, val_loader, args, log_dir, writer, logger=None, timer=None,
               is_main=True):
    timer = timer or Timer()
    total_down = total_up = 0.0
    n_epochs = math.ceil(args.num_epochs)
    for epoch in range(n_epochs):
        frac = args.num_epochs - epoch if epoch == n_epochs - 1 else 1
        _, down, up = run_batches(model, opt, scheduler, train_loader, args, timer, training=True, epoch=epoch,
                                  epoch_fraction=frac, logger=logger, writer=writer, is_main=is_main)
        total_down += down.sum().item() / 2 ** 20
        total_up += up.sum().item() / 2 ** 20
    if is_main:
        nc = train_loader.dataset.num_clients
        print("Total Download (MiB): {:0.2f}".format(total_down))
        print("Total Upload (MiB): {:0.2f}".format(total_up))
        print("Avg Download Per Client: {:0.2f}".format(total_down / nc))
        print("Avg Upload Per Client: {:0.2f}".format(total_up / nc))
    model.save_pretrained(log_dir)
    return test_gpt2(model, val_loader, args, timer=timer, writer=writer, is_main=is_main)


# ---------------------------------------------------------------- data / model construction
def get_data_loaders(args, tokenizer, rank=0):
    if getattr(args, "synthetic", False):
        if rank == 0:
            data_utils.write_synthetic_persona(args.dataset_dir, num_clients=args.num_clients or 64,
                                               num_candidates=max(2, args.num_candidates), seed=args.seed)
        init_runtime(args).barrier()
    common = dict(tokenizer=tokenizer, num_candidates=args.num_candidates, max_history=args.max_history,
                  dataset_dir=args.dataset_dir, dataset_name=args.dataset_name)
    train_ds = FedPERSONA(personality_permutations=args.personality_permutations, do_iid=args.do_iid,
                          num_clients=args.num_clients if args.do_iid else None, train=True, **common)
    val_ds = FedPERSONA(personality_permutations=1, train=False, **common)
    sampler = FedSampler(train_ds, args.num_workers, args.local_batch_size, seed=args.seed)
    train_loader = DataLoader(train_ds, batch_sampler=sampler, collate_fn=personachat_collate_fn,
                              num_workers=args.train_dataloader_workers)
    from .data_utils import ShardedEvalSampler
    val_sampler = ShardedEvalSampler(len(val_ds), args.valid_batch_size * args.num_workers, args.valid_batch_size)
    val_loader = DataLoader(val_ds, batch_sampler=val_sampler, collate_fn=personachat_collate_fn,
                            num_workers=args.val_dataloader_workers)
    return train_loader, val_loader


def build_model_and_tokenizer(args):
    ckpt = args.finetune_path if (args.do_finetune and not args.do_test) else args.model_checkpoint
    cls = GPT2DoubleHeadsModel if "gpt2" in args.model_checkpoint else OpenAIGPTDoubleHeadsModel
    if os.path.isdir(str(ckpt)):
        tokenizer = load_tokenizer(ckpt)           # GPT-2 BPE files when the directory has them
        model = cls.from_pretrained(ckpt)
    elif args.model_checkpoint == "gpt2-tiny":
        tokenizer = SimpleTokenizer(TINY["vocab_size"])
        model = cls(GPT2Config(**TINY))
    else:
        model = cls()
        tokenizer = SimpleTokenizer(model.config.vocab_size)
    return model, tokenizer


def train(argv=None):
    args = parse_args(default_lr=4e-2, argv=argv)
    if not args.dataset_name:
        args.dataset_name = "PERSONA"
    args.num_results_train = 1          # the train callback returns the loss only
    rt = init_runtime(args)
    is_main = rt.rank == 0
    if is_main:
        print(args)
    timer = Timer()
    np.random.seed(args.seed)
    torch.random.manual_seed(args.seed)

    model, tokenizer = build_model_and_tokenizer(args)
    args.len_tokenizer = len(tokenizer)
    log_dir = make_logdir(args)
    writer = None
    if is_main:
        os.makedirs(log_dir, exist_ok=True)
        if args.use_tensorboard:
            from torch.utils.tensorboard import SummaryWriter
            writer = SummaryWriter(log_dir=log_dir)
        model.config.to_json_file(os.path.join(log_dir, "config.json"))
    add_special_tokens_(model, tokenizer)
    if is_main:
        tokenizer.save_pretrained(log_dir)
    optimizer = SGD(model.parameters(), lr=1)
    train_loader, val_loader = get_data_loaders(args, tokenizer, rt.rank)
    if args.num_clients is None:
        args.num_clients = train_loader.dataset.num_clients

    fed_model = FedModel(model, compute_loss_train, args, compute_loss_val)
    fed_model.attach_sampler(train_loader.batch_sampler)      # a rank's loader tokenises only its own clients
    fed_model.attach_val_sampler(val_loader.batch_sampler)
    fed_opt = FedOptimizer(optimizer, args)
    if args.resume:
        fed_model.load_fedstate(args.resume, fed_opt)
    spe = steps_per_epoch(args.local_batch_size, train_loader.dataset, args.num_workers)
    schedule = PiecewiseLinear([0, args.num_epochs * spe], [args.lr_scale, 0.0])
    scheduler = LambdaLR(fed_opt, lr_lambda=lambda step: schedule(step))
    if args.resume:
        fed_model.restore_scheduler(scheduler)
    if is_main:
        print("Steps per epoch", spe)
    if args.do_finetune:
        out = test_gpt2(fed_model, val_loader, args, logger=TableLogger(), timer=timer, writer=writer, is_main=is_main)
    else:
        out = train_gpt2(fed_model, fed_opt, scheduler, train_loader, val_loader, args, log_dir, writer,
                         logger=TableLogger(), timer=timer, is_main=is_main)
    fed_model.finalize()
    if args.do_checkpoint:
        os.makedirs(args.checkpoint_path, exist_ok=True)
        fed_model.save_fedstate(os.path.join(args.checkpoint_path, "gpt2.fedstate.pt"), fed_opt, scheduler)
    rt.shutdown()
    return out


if __name__ == "__main__":
    train()
