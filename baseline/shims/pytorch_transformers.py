"""Stand-in for ``pytorch_transformers`` (absent on this image) built on the installed ``transformers``
package: the names the reference imports (gpt2_train.py:4-6, data_utils/fed_persona.py:15) with the
old calling convention -- ``forward(input_ids, token_type_ids=, mc_token_ids=, lm_labels=, mc_labels=)``
returning ``(lm_loss, mc_loss, lm_logits, mc_logits)``, ``lm_labels`` using -1 for "ignore" -- mapped onto
HuggingFace's ``GPT2DoubleHeadsModel`` (``labels`` / -100).  Random initialisation only: there is no network
for pretrained weights.  Baseline-only code; nothing of commefficient_b200 is imported."""
import os

import _ref_compat  # noqa: F401

CONFIG_NAME = "config.json"
WEIGHTS_NAME = "pytorch_model.bin"


def cached_path(url_or_filename, cache_dir=None):
    raise RuntimeError("no network: cannot fetch %s" % (url_or_filename,))


class _DoubleHeadsAdapter:
    """Mixin giving a HuggingFace double-heads model the pytorch_transformers call signature."""

    def forward(self, input_ids=None, token_type_ids=None, mc_token_ids=None, lm_labels=None, mc_labels=None, **kw):
        labels = None
        if lm_labels is not None:
            labels = lm_labels.masked_fill(lm_labels == -1, -100)
        out = super().forward(input_ids=input_ids, token_type_ids=token_type_ids, mc_token_ids=mc_token_ids,
                              labels=labels, mc_labels=mc_labels, return_dict=True, use_cache=False)
        res = []
        if lm_labels is not None:
            res.append(out.loss)
        if mc_labels is not None:
            res.append(out.mc_loss)
        res.extend([out.logits, out.mc_logits])
        return tuple(res)


GPT2DoubleHeadsModel = OpenAIGPTDoubleHeadsModel = None
if os.environ.get("CEB_REF_GPT2") == "1":
    # Built at import time (and only for GPT-2 runs: importing ``transformers`` takes seconds) as module-level
    # classes, because the reference hands the model to its worker processes by pickling it (spawn).
    import transformers as _tr

    class GPT2DoubleHeadsModel(_DoubleHeadsAdapter, _tr.GPT2DoubleHeadsModel):
        @classmethod
        def from_pretrained(cls, name_or_path, *a, **kw):
            return cls(_tr.GPT2Config())          # random-init GPT-2 small: no network for pretrained weights

    class OpenAIGPTDoubleHeadsModel(_DoubleHeadsAdapter, _tr.OpenAIGPTDoubleHeadsModel):
        @classmethod
        def from_pretrained(cls, name_or_path, *a, **kw):
            return cls(_tr.OpenAIGPTConfig())


class _Tokenizer:
    """Minimal tokenizer object (synthetic token ids are generated directly by the harness)."""

    def __init__(self, vocab=50257):
        self.vocab = vocab
        self.added = []

    @classmethod
    def from_pretrained(cls, *a, **kw):
        return cls()

    def __len__(self):
        return self.vocab + len(self.added)

    def add_special_tokens(self, mapping):
        n = 0
        for v in mapping.values():
            for tok in (v if isinstance(v, (list, tuple)) else [v]):
                if tok not in self.added:
                    self.added.append(tok)
                    n += 1
        return n

    def save_pretrained(self, path):
        pass


GPT2Tokenizer = OpenAIGPTTokenizer = _Tokenizer


def AdamW(*a, **kw):
    import torch
    return torch.optim.AdamW(*a, **kw)
