"""GPT-2 with a language-modelling head and a multiple-choice head
("double heads"), the model the reference fine-tunes on PersonaChat through the
external ``pytorch_transformers`` package
(/root/reference/CommEfficient/gpt2_train.py:4-6,77-112,262-285).

This is a self-contained implementation: the package is not installed here and
the installed ``transformers`` uses different keyword names.
Kept compatible:
  * call signature ``model(input_ids, token_type_ids=, mc_token_ids=, lm_labels=,
    mc_labels=)`` and tuple return ``(lm_loss, mc_loss, lm_logits, mc_logits)``
    (losses present only when the labels are given);
  * LM loss = shifted cross-entropy with ``ignore_index=-1``; MC loss =
    cross-entropy over the candidates of ``Linear(n_embd -> 1)`` applied to the
    hidden state at ``mc_token_ids``;
  * Hugging Face ``state_dict`` key names and ``Conv1D`` (in, out) weight layout, so
    ``save_pretrained`` output loads into HF GPT2DoubleHeadsModel and vice versa;
  * ``resize_token_embeddings`` (+5 special tokens => d = 124 444 417 for GPT-2 small).
"""
from __future__ import annotations

import json
import math
import os
from dataclasses import asdict, dataclass
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import nn as nnops

__all__ = ["GPT2Config", "GPT2DoubleHeadsModel", "OpenAIGPTDoubleHeadsModel"]

CONFIG_NAME = "config.json"
WEIGHTS_NAME = "pytorch_model.bin"


@dataclass
class GPT2Config:
    vocab_size: int = 50257
    n_positions: int = 1024
    n_ctx: int = 1024
    n_embd: int = 768
    n_layer: int = 12
    n_head: int = 12
    resid_pdrop: float = 0.1
    embd_pdrop: float = 0.1
    attn_pdrop: float = 0.1
    layer_norm_epsilon: float = 1e-5
    initializer_range: float = 0.02
    summary_first_dropout: float = 0.1
    post_ln: bool = False          # GPT-1 block order: x = ln_1(x + attn(x)); x = ln_2(x + mlp(x)); no final LayerNorm

    def to_json_file(self, path):
        with open(path, "w") as f:
            json.dump(asdict(self), f, indent=2)

    @classmethod
    def from_json_file(cls, path):
        with open(path) as f:
            raw = json.load(f)
        known = {k: v for k, v in raw.items() if k in cls.__dataclass_fields__}
        return cls(**known)


class Conv1D(nn.Module):
    """Linear layer with the (in_features, out_features) weight layout of GPT-2."""

    def __init__(self, nf, nx, std=0.02):
        super().__init__()
        self.nf = nf
        self.weight = nn.Parameter(torch.empty(nx, nf).normal_(0.0, std))
        self.bias = nn.Parameter(torch.zeros(nf))

    def forward(self, x):
        shape = x.shape[:-1] + (self.nf,)
        y = nnops.linear(x.reshape(-1, x.shape[-1]), self.weight.t(), self.bias)
        return y.view(shape)


class Attention(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.n_head = cfg.n_head
        self.c_attn = Conv1D(3 * cfg.n_embd, cfg.n_embd, cfg.initializer_range)
        self.c_proj = Conv1D(cfg.n_embd, cfg.n_embd, cfg.initializer_range)
        self.attn_pdrop = cfg.attn_pdrop
        self.resid_dropout = nn.Dropout(cfg.resid_pdrop)

    def forward(self, x, salt=0, fuse_resid_dropout=False):
        """``fuse_resid_dropout``: the caller applies the residual dropout inside the add + LayerNorm kernel
        that consumes this branch, so it is skipped here."""
        B, T, E = x.shape
        y = nnops.attention(self.c_attn(x), B, T, self.n_head,
                            p_drop=self.attn_pdrop if self.training else 0.0, salt=salt)
        y = self.c_proj(y)
        return y if fuse_resid_dropout else self.resid_dropout(y)


class MLP(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.c_fc = Conv1D(4 * cfg.n_embd, cfg.n_embd, cfg.initializer_range)
        self.c_proj = Conv1D(cfg.n_embd, 4 * cfg.n_embd, cfg.initializer_range)
        self.dropout = nn.Dropout(cfg.resid_pdrop)

    def forward(self, x, fuse_resid_dropout=False):
        fc, pr = self.c_fc, self.c_proj
        y = nnops.mlp_gelu(x.reshape(-1, x.shape[-1]), fc.weight.t(), fc.bias, pr.weight.t(), pr.bias)
        y = y.view(x.shape[:-1] + (pr.nf,))
        return y if fuse_resid_dropout else self.dropout(y)


class Block(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.ln_1 = nn.LayerNorm(cfg.n_embd, eps=cfg.layer_norm_epsilon)
        self.attn = Attention(cfg)
        self.ln_2 = nn.LayerNorm(cfg.n_embd, eps=cfg.layer_norm_epsilon)
        self.mlp = MLP(cfg)

    def forward(self, x, post_ln=False):
        ln1, ln2 = self.ln_1, self.ln_2
        if post_ln:                                                   # GPT-1 order
            x = nnops.layer_norm(x + self.attn(x), ln1.weight, ln1.bias, ln1.eps)
            return nnops.layer_norm(x + self.mlp(x), ln2.weight, ln2.bias, ln2.eps)
        x = x + self.attn(nnops.layer_norm(x, ln1.weight, ln1.bias, ln1.eps))
        return x + self.mlp(nnops.layer_norm(x, ln2.weight, ln2.bias, ln2.eps))


class GPT2Model(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.wte = nn.Embedding(cfg.vocab_size, cfg.n_embd)
        self.wpe = nn.Embedding(cfg.n_positions, cfg.n_embd)
        self.drop = nn.Dropout(cfg.embd_pdrop)
        self.h = nn.ModuleList([Block(cfg) for _ in range(cfg.n_layer)])
        self.post_ln = bool(getattr(cfg, "post_ln", False))
        if not self.post_ln:
            self.ln_f = nn.LayerNorm(cfg.n_embd, eps=cfg.layer_norm_epsilon)
        nn.init.normal_(self.wte.weight, 0.0, cfg.initializer_range)
        nn.init.normal_(self.wpe.weight, 0.0, cfg.initializer_range)

    def forward(self, input_ids, token_type_ids=None, flat_params=None):
        # native path: every dropout of the trunk is applied INSIDE the kernel that produces / consumes the
        # tensor (embedding gather, attention probabilities, the residual add + LayerNorm); masks come from a
        # counter hash of (device seed, site salt, position) and are regenerated in the backward
        fused = self.training and nnops.fused_dropout_available(self.wte.weight)
        if fused:
            nnops.advance_dropout(input_ids.device)
        x = nnops.embed(self.wte.weight, self.wpe.weight, input_ids, token_type_ids,
                        p_drop=self.drop.p if self.training else 0.0, salt=1)
        # Same math as ``for block in self.h: x = block(x)`` followed by ln_f, regrouped so that every
        # residual add is fused with the LayerNorm that consumes its result (ops/nn.py:add_layer_norm):
        # x + attn(ln_1(x)) feeds ln_2 of the same block, x + mlp(ln_2(x)) feeds ln_1 of the next block
        # (or ln_f after the last one).
        blocks = list(self.h)
        if self.post_ln:
            # GPT-1 (OpenAIGPTDoubleHeadsModel): the LayerNorm FOLLOWS each residual add and its output is the
            # stream; the same fused add + LayerNorm kernel serves, its normalised output is what is carried on
            for i, block in enumerate(blocks):
                pr = block.attn.resid_dropout.p if fused else 0.0
                _, x = nnops.add_layer_norm(x, block.attn(x, salt=16 * i + 2, fuse_resid_dropout=fused),
                                            block.ln_1.weight, block.ln_1.bias, block.ln_1.eps, p_drop=pr, salt=16 * i + 3)
                pr = block.mlp.dropout.p if fused else 0.0
                _, h = nnops.add_layer_norm(x.float(), block.mlp(x, fuse_resid_dropout=fused), block.ln_2.weight,
                                            block.ln_2.bias, block.ln_2.eps, p_drop=pr, salt=16 * i + 4)
                x = h.float() if i + 1 < len(blocks) else h
            return x
        first = blocks[0].ln_1 if blocks else self.ln_f
        h = nnops.layer_norm(x, first.weight, first.bias, first.eps)
        report = (flat_params is not None and flat_params.on_range_final is not None and torch.is_grad_enabled()
                  and x.requires_grad)
        for i, block in enumerate(blocks):
            if report:
                # The gradient of the stream entering block i is complete only after the whole backward of block i
                # (and of the LayerNorm that follows it) has run: everything from this block's attention weights
                # (from its ln_1 for block 0) up to and including the next block's ln_1 / ln_f is final then, and
                # the consumer may start sketching it while blocks i-1 .. 0 are still in their backward pass.
                first = block.ln_1.weight if i == 0 else block.attn.c_attn.weight
                nxt_ln = blocks[i + 1].ln_1 if i + 1 < len(blocks) else self.ln_f
                params = [p for p in flat_params.params]
                ids = {id(p): k for k, p in enumerate(params)}
                lo_k, hi_k = ids.get(id(first)), ids.get(id(nxt_ln.bias))
                if lo_k is not None and hi_k is not None and hi_k > lo_k:
                    lo = flat_params.offsets[lo_k]
                    hi = flat_params.offsets[hi_k] + flat_params.sizes[hi_k]
                    x.register_hook(lambda g, lo=lo, hi=hi: flat_params.range_final(lo, hi))
            pr = block.attn.resid_dropout.p if fused else 0.0
            x, h = nnops.add_layer_norm(x, block.attn(h, salt=16 * i + 2, fuse_resid_dropout=fused), block.ln_2.weight,
                                        block.ln_2.bias, block.ln_2.eps, p_drop=pr, salt=16 * i + 3)
            nxt = blocks[i + 1].ln_1 if i + 1 < len(blocks) else self.ln_f
            pr = block.mlp.dropout.p if fused else 0.0
            x, h = nnops.add_layer_norm(x, block.mlp(h, fuse_resid_dropout=fused), nxt.weight, nxt.bias, nxt.eps,
                                        p_drop=pr, salt=16 * i + 4)
        return h


class SequenceSummary(nn.Module):
    """Hidden state at a given token -> dropout -> Linear(n_embd, 1)."""

    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.first_dropout = nn.Dropout(cfg.summary_first_dropout)
        self.summary = nn.Linear(cfg.n_embd, 1)
        nn.init.normal_(self.summary.weight, 0.0, cfg.initializer_range)
        nn.init.zeros_(self.summary.bias)

    def forward(self, hidden, cls_index):
        # hidden (N, T, E), cls_index (N,)
        gathered = hidden[torch.arange(hidden.shape[0], device=hidden.device), cls_index]
        return self.summary(self.first_dropout(gathered)).squeeze(-1)


class GPT2DoubleHeadsModel(nn.Module):
    def __init__(self, config: Optional[GPT2Config] = None):
        super().__init__()
        self.config = config or GPT2Config()
        self.transformer = GPT2Model(self.config)
        self.lm_head = nn.Linear(self.config.n_embd, self.config.vocab_size, bias=False)
        self.multiple_choice_head = SequenceSummary(self.config)
        self.tie_weights()

    # -- HF-style utilities --------------------------------------------------------
    def tie_weights(self):
        self.lm_head.weight = self.transformer.wte.weight

    def resize_token_embeddings(self, new_num_tokens: int):
        old = self.transformer.wte
        if new_num_tokens == old.num_embeddings:
            return old
        new = nn.Embedding(new_num_tokens, old.embedding_dim).to(old.weight.device)
        nn.init.normal_(new.weight, 0.0, self.config.initializer_range)
        n = min(old.num_embeddings, new_num_tokens)
        with torch.no_grad():
            new.weight[:n] = old.weight[:n]
        self.transformer.wte = new
        self.lm_head = nn.Linear(old.embedding_dim, new_num_tokens, bias=False)
        self.config.vocab_size = new_num_tokens
        self.tie_weights()
        return new

    @classmethod
    def from_pretrained(cls, path_or_name, **kw):
        """Load a checkpoint directory in the published layout: ``config.json`` plus ``pytorch_model.bin`` or
        ``model.safetensors``, with or without the ``transformer.`` key prefix (``GPT2Model`` checkpoints, such as
        the original ``gpt2`` files, have none).  Heads the checkpoint lacks keep their fresh initialisation.  Any
        other argument (e.g. the hub name ``gpt2`` -- there is no network) builds a randomly initialised model."""
        if isinstance(path_or_name, str) and os.path.isdir(path_or_name):
            cfg = GPT2Config.from_json_file(os.path.join(path_or_name, CONFIG_NAME))
            model = cls(cfg)
            bin_fn = os.path.join(path_or_name, WEIGHTS_NAME)
            st_fn = os.path.join(path_or_name, "model.safetensors")
            if os.path.exists(bin_fn):
                state = torch.load(bin_fn, map_location="cpu")
            elif os.path.exists(st_fn):
                from safetensors.torch import load_file
                state = load_file(st_fn, device="cpu")
            else:
                raise FileNotFoundError("no %s or model.safetensors in %s" % (WEIGHTS_NAME, path_or_name))
            state = {k: v for k, v in state.items() if not k.endswith(".attn.bias")
                     and not k.endswith(".attn.masked_bias")}
            # GPT-1 checkpoints name the two embedding tables differently
            state = {k.replace("tokens_embed.", "wte.").replace("positions_embed.", "wpe."): v for k, v in state.items()}
            if not any(k.startswith("transformer.") for k in state):
                state = {("transformer." + k if k.split(".")[0] in ("wte", "wpe", "h", "ln_f") else k): v
                         for k, v in state.items()}
            if "lm_head.weight" not in state and "transformer.wte.weight" in state:
                state["lm_head.weight"] = state["transformer.wte.weight"]          # tied
            own = model.state_dict()
            trunk = [k for k in own if k.startswith("transformer.")]
            missing = [k for k in trunk if k not in state]
            if missing:
                raise KeyError("checkpoint %s lacks %d trunk tensors, e.g. %s" % (path_or_name, len(missing), missing[:3]))
            bad = [k for k in own if k in state and tuple(state[k].shape) != tuple(own[k].shape)]
            if bad:
                raise ValueError("checkpoint %s: shape mismatch for %s" % (path_or_name, bad[:3]))
            model.load_state_dict(state, strict=False)
            model.tie_weights()
            return model
        return cls(GPT2Config(**kw))

    def save_pretrained(self, save_directory):
        os.makedirs(save_directory, exist_ok=True)
        self.config.to_json_file(os.path.join(save_directory, CONFIG_NAME))
        torch.save(self.state_dict(), os.path.join(save_directory, WEIGHTS_NAME))

    @staticmethod
    def lm_loss(lm_logits, lm_labels):
        """Mean shifted cross-entropy over the non-ignored target tokens (ignore_index -1)."""
        rows, mask = shifted_lm_rows(lm_logits, lm_labels)
        return rows.sum() / mask.sum().clamp_min(1)

    def lm_rows(self, hidden, lm_labels):
        """Per-position shifted LM cross-entropy ``[N, T]`` and its mask straight from the hidden
        states: the LM head only runs for positions that carry a label (``ops/nn.py``)."""
        T = lm_labels.shape[-1]
        shifted = _shift_labels(lm_labels)
        rows = nnops.lm_head_cross_entropy(hidden.reshape(-1, hidden.shape[-1]), self.lm_head.weight,
                                           shifted.reshape(-1), ignore_index=-1)
        return rows.view(-1, T), (shifted != -1)

    # -- forward -------------------------------------------------------------------
    def forward(self, input_ids, token_type_ids=None, mc_token_ids=None, lm_labels=None, mc_labels=None,
                need_logits=True, return_hidden=False, **_ignored):
        """Outputs follow pytorch_transformers: ``([lm_loss], [mc_loss], lm_logits, mc_logits)``.
        ``need_logits=False`` leaves ``lm_logits`` as None and computes the LM loss through
        ``lm_rows`` (no ``[N*C*T, V]`` logits for unlabeled positions); ``return_hidden=True``
        appends the final hidden states."""
        shape = input_ids.shape                      # (B, C, T)
        T = shape[-1]
        flat_ids = input_ids.reshape(-1, T)
        flat_types = token_type_ids.reshape(-1, T) if token_type_ids is not None else None
        hidden = self.transformer(flat_ids, flat_types, flat_params=getattr(self, "_flat_params", None))   # (B*C, T, E)
        lm_logits = None
        if need_logits:
            lm_logits = nnops.linear(hidden.reshape(-1, hidden.shape[-1]), self.lm_head.weight, None)
            lm_logits = lm_logits.view(*shape, -1)
        outputs = [lm_logits]
        mc_logits = None
        if mc_token_ids is not None:
            mc_logits = self.multiple_choice_head(hidden, mc_token_ids.reshape(-1)).view(shape[:-1])
        outputs.append(mc_logits)
        if mc_labels is not None:
            mc_loss = F.cross_entropy(mc_logits.float().view(-1, mc_logits.shape[-1]), mc_labels.view(-1))
            outputs.insert(0, mc_loss)
        if lm_labels is not None:
            if lm_logits is not None:
                outputs.insert(0, self.lm_loss(lm_logits, lm_labels))
            else:
                rows, mask = self.lm_rows(hidden, lm_labels)
                outputs.insert(0, rows.sum() / mask.sum().clamp_min(1))
        if return_hidden:
            outputs.append(hidden)
        return tuple(outputs)


def _shift_labels(lm_labels):
    """Position t is scored against label t+1; the last position is ignored."""
    T = lm_labels.shape[-1]
    flat = lm_labels.reshape(-1, T)
    return torch.cat([flat[:, 1:], flat.new_full((flat.shape[0], 1), -1)], dim=1)


def shifted_lm_rows(lm_logits, lm_labels):
    """Per-position LM cross-entropy ``[N, T]`` (0 at ignored positions) and the matching mask.
    Position t is scored against label t+1; instead of slicing the logits (which would copy the
    whole ``[N, T, V]`` tensor) the LABELS are shifted and the last position is ignored."""
    T = lm_labels.shape[-1]
    shifted = _shift_labels(lm_labels)
    logits2d = lm_logits.reshape(-1, lm_logits.shape[-1])
    rows = nnops.cross_entropy_rows(logits2d, shifted.reshape(-1), ignore_index=-1)
    return rows.view(-1, T), (shifted != -1)


class OpenAIGPTDoubleHeadsModel(GPT2DoubleHeadsModel):
    """The reference falls back to the GPT-1 class when ``--model_checkpoint`` does not
    contain "gpt2" (gpt2_train.py:263-268).  Same double-head interface; GPT-1's architecture: 40 478 BPE
    tokens, 512 positions, LayerNorm AFTER each residual add (``post_ln``) and no final LayerNorm."""

    def __init__(self, config: Optional[GPT2Config] = None):
        config = config or GPT2Config(vocab_size=40478, n_positions=512, n_ctx=512, post_ln=True)
        config.post_ln = True
        super().__init__(config)
