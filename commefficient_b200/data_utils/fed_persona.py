"""PersonaChat as a federated dataset: one client per personality, one item per
utterance, each item a set of candidate replies rendered as token sequences.

Formats and tensor layout follow
/root/reference/CommEfficient/data_utils/fed_persona.py:31-393:
  files   ``client{i}.json`` (list of dialogs), ``validation.json``, ``stats.json`` with
          ``dialogs_per_client`` / ``train_utterances_per_dialog`` / ``val_utterances_per_dialog``
  item    ``(client_id, input_ids[C][T], mc_token_ids[C], lm_labels[C][T], mc_labels,
          token_type_ids[C][T])``; the last candidate is the true reply
  batch   ``personachat_collate_fn`` pads to the round's longest sequence (0, or -1 for
          ``lm_labels``) and reshapes to (B, C, T).
Departures: a client's json is cached after first use instead of being
re-read from disk for every sample; validation indexing is re-derived (Q15);
``personality_permutations`` > 1 yields that many permuted copies of the item's
candidates instead of silently keeping only the last permutation (Q14).
The raw corpus cannot be downloaded here; ``data_utils.synthetic.write_synthetic_persona``
writes templated dialogs in the same format, and ``SimpleTokenizer`` stands in for
the GPT-2 BPE tokenizer (whose vocabulary files are not on this machine).
"""
from __future__ import annotations

import json
import os
import random
from collections import defaultdict
from functools import lru_cache
from itertools import chain

import numpy as np
import torch
from torch.nn.utils.rnn import pad_sequence

from .fed_dataset import FedDataset

__all__ = ["FedPERSONA", "personachat_collate_fn", "SimpleTokenizer", "SPECIAL_TOKENS", "MODEL_INPUTS"]

SPECIAL_TOKENS = ["<bos>", "<eos>", "<speaker1>", "<speaker2>", "<pad>"]
MODEL_INPUTS = ["input_ids", "mc_token_ids", "lm_labels", "mc_labels", "token_type_ids"]
PADDED_INPUTS = ["input_ids", "lm_labels", "token_type_ids"]
ATTR_TO_SPECIAL_TOKEN = {"bos_token": "<bos>", "eos_token": "<eos>", "pad_token": "<pad>",
                         "additional_special_tokens": ("<speaker1>", "<speaker2>")}


class SimpleTokenizer:
    """Whitespace tokenizer with a hashed fixed-size vocabulary -- API subset of the HF
    GPT-2 tokenizer that the data pipeline and driver use (``tokenize``,
    ``convert_tokens_to_ids``, ``add_special_tokens``, ``encoder``, ``__len__``,
    ``save_pretrained`` / ``from_pretrained``)."""

    def __init__(self, vocab_size=50257):
        self.vocab_size = int(vocab_size)
        self.encoder = _RangeDict(self.vocab_size)
        self.special = {}

    @classmethod
    def from_pretrained(cls, path_or_name, **kw):
        fn = os.path.join(str(path_or_name), "simple_tokenizer.json")
        if os.path.exists(fn):
            with open(fn) as f:
                blob = json.load(f)
            tok = cls(blob["vocab_size"])
            tok.special = dict(blob["special"])
            return tok
        return cls(**kw)

    def save_pretrained(self, directory):
        os.makedirs(directory, exist_ok=True)
        with open(os.path.join(directory, "simple_tokenizer.json"), "w") as f:
            json.dump({"vocab_size": self.vocab_size, "special": self.special}, f)

    def __len__(self):
        return self.vocab_size + len(self.special)

    def add_special_tokens(self, mapping):
        added = 0
        for value in mapping.values():
            for tok in ([value] if isinstance(value, str) else value):
                if tok not in self.special:
                    self.special[tok] = self.vocab_size + len(self.special)
                    added += 1
        return added

    def tokenize(self, text):
        return text.lower().split()

    def _id(self, token):
        if token in self.special:
            return self.special[token]
        h = 2166136261
        for ch in token.encode("utf-8"):
            h = ((h ^ ch) * 16777619) & 0xFFFFFFFF
        return h % self.vocab_size

    def convert_tokens_to_ids(self, tokens):
        if isinstance(tokens, str):
            return self._id(tokens)
        return [self._id(t) for t in tokens]


def _bytes_to_unicode():
    """GPT-2's reversible byte -> printable-unicode table: the 188 printable latin-1 bytes map to themselves, the
    other 68 to code points 256, 257, ... (so that no byte becomes whitespace / a control character in the vocabulary)."""
    keep = list(range(ord("!"), ord("~") + 1)) + list(range(0xA1, 0xAD)) + list(range(0xAE, 0x100))
    table, extra = {}, 0
    for b in range(256):
        if b in keep:
            table[b] = chr(b)
        else:
            table[b] = chr(256 + extra)
            extra += 1
    return table


class BPETokenizer:
    """Byte-level BPE tokenizer reading GPT-2's published files (``vocab.json`` + ``merges.txt``): the tokenizer the
    reference gets from ``pytorch_transformers.GPT2Tokenizer`` (gpt2_train.py:259-262), restated without that
    dependency.  Same surface as ``SimpleTokenizer``.  Used whenever ``--model_checkpoint`` / ``--finetune_path`` is a
    directory holding the two files (they cannot be downloaded on this machine; ``tests/test_cpu_units.py`` checks
    this implementation against ``transformers``' on a generated vocabulary)."""

    VOCAB, MERGES, SPECIAL = "vocab.json", "merges.txt", "added_tokens.json"

    def __init__(self, vocab_file, merges_file, special=None):
        import regex
        with open(vocab_file, encoding="utf-8") as f:
            self.encoder = json.load(f)
        with open(merges_file, encoding="utf-8") as f:
            lines = [ln.rstrip("\n") for ln in f]
        merges = [tuple(ln.split()) for ln in lines if ln and not ln.startswith("#version") and len(ln.split()) == 2]
        self.ranks = {pair: i for i, pair in enumerate(merges)}
        self.byte_table = _bytes_to_unicode()
        self.special = dict(special or {})
        self.pattern = regex.compile(
            r"'s|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+")
        self._cache = {}
        self._merges_text = "\n".join(lines)

    # -- persistence ---------------------------------------------------------------------
    @classmethod
    def has_files(cls, directory) -> bool:
        d = str(directory)
        return os.path.exists(os.path.join(d, cls.VOCAB)) and os.path.exists(os.path.join(d, cls.MERGES))

    @classmethod
    def from_pretrained(cls, directory, **_kw):
        d = str(directory)
        special = {}
        fn = os.path.join(d, cls.SPECIAL)
        if os.path.exists(fn):
            with open(fn, encoding="utf-8") as f:
                special = json.load(f)
        return cls(os.path.join(d, cls.VOCAB), os.path.join(d, cls.MERGES), special)

    def save_pretrained(self, directory):
        os.makedirs(directory, exist_ok=True)
        with open(os.path.join(directory, self.VOCAB), "w", encoding="utf-8") as f:
            json.dump(self.encoder, f, ensure_ascii=False)
        with open(os.path.join(directory, self.MERGES), "w", encoding="utf-8") as f:
            f.write(self._merges_text + "\n")
        with open(os.path.join(directory, self.SPECIAL), "w", encoding="utf-8") as f:
            json.dump(self.special, f)

    def __len__(self):
        return len(self.encoder) + len(self.special)

    def add_special_tokens(self, mapping):
        added = 0
        for value in mapping.values():
            for tok in ([value] if isinstance(value, str) else value):
                if tok not in self.special and tok not in self.encoder:
                    self.special[tok] = len(self.encoder) + len(self.special)
                    added += 1
        return added

    # -- BPE -------------------------------------------------------------------------------
    def _bpe(self, word: str):
        cached = self._cache.get(word)
        if cached is not None:
            return cached
        parts = list(word)
        while len(parts) > 1:
            best, best_rank = None, None
            for i in range(len(parts) - 1):
                r = self.ranks.get((parts[i], parts[i + 1]))
                if r is not None and (best_rank is None or r < best_rank):
                    best, best_rank = (parts[i], parts[i + 1]), r
            if best is None:
                break
            merged, i = [], 0
            while i < len(parts):
                if i + 1 < len(parts) and (parts[i], parts[i + 1]) == best:
                    merged.append(parts[i] + parts[i + 1])
                    i += 2
                else:
                    merged.append(parts[i])
                    i += 1
            parts = merged
        self._cache[word] = parts
        return parts

    def tokenize(self, text):
        out = []
        for piece in self.pattern.findall(text):
            out.extend(self._bpe("".join(self.byte_table[b] for b in piece.encode("utf-8"))))
        return out

    def _id(self, token):
        if token in self.special:
            return self.special[token]
        return self.encoder[token]

    def convert_tokens_to_ids(self, tokens):
        if isinstance(tokens, str):
            return self._id(tokens)
        return [self._id(t) for t in tokens]


class _RangeDict:
    """len()-only stand-in for the HF ``encoder`` dict (avoids a 50k-entry dict)."""

    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n


def load_tokenizer(directory):
    """The tokenizer saved in ``directory``: GPT-2 BPE files if present, else the hashed whitespace stand-in."""
    if BPETokenizer.has_files(directory):
        return BPETokenizer.from_pretrained(directory)
    return SimpleTokenizer.from_pretrained(directory)


def tokenize(obj, tokenizer):
    if isinstance(obj, str):
        return tokenizer.convert_tokens_to_ids(tokenizer.tokenize(obj))
    if isinstance(obj, dict):
        return {k: tokenize(v, tokenizer) for k, v in obj.items()}
    return [tokenize(o, tokenizer) for o in obj]


def build_input_from_segments(persona, history, reply, tokenizer, lm_labels=False, with_eos=True):
    """[bos]+persona | turns prefixed with alternating speaker tokens | reply+[eos]."""
    bos, eos, spk1, spk2 = tokenizer.convert_tokens_to_ids(SPECIAL_TOKENS[:-1])
    segments = [[bos] + list(chain.from_iterable(persona))] + list(history)
    segments.append(list(reply) + ([eos] if with_eos else []))
    n = len(segments)
    tagged = [segments[0]]
    for i, seg in enumerate(segments[1:]):
        tagged.append([spk2 if (n - i) % 2 == 0 else spk1] + seg)
    ids = list(chain.from_iterable(tagged))
    types = [spk2 if i % 2 else spk1 for i, seg in enumerate(tagged) for _ in seg]
    labels = [-1] * len(ids)
    if lm_labels:
        prefix = sum(len(s) for s in tagged[:-1])
        labels = [-1] * prefix + [-1] + tagged[-1][1:]
    return {"input_ids": ids, "token_type_ids": types, "mc_token_ids": len(ids) - 1, "lm_labels": labels}


def raw_to_input(tokenizer, personality, history, candidates):
    personality = tokenize(personality, tokenizer)
    history = tokenize(history, tokenizer)
    candidates = tokenize(candidates, tokenizer)
    cols = defaultdict(list)
    last = len(candidates) - 1
    for j, cand in enumerate(candidates):
        inst = build_input_from_segments(personality, history, cand, tokenizer, lm_labels=(j == last))
        for name, arr in inst.items():
            cols[name].append(arr)
    out = []
    for name in MODEL_INPUTS:
        if name == "mc_labels":
            out.append(last)
        elif name == "mc_token_ids":
            out.append(torch.tensor(cols[name]))
        else:
            out.append([torch.tensor(a) for a in cols[name]])
    return tuple(out)


class FedPERSONA(FedDataset):
    def __init__(self, tokenizer, num_candidates, max_history, personality_permutations, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.tokenizer = tokenizer
        self.num_candidates = num_candidates
        self.max_history = max_history
        self.personality_permutations = personality_permutations
        if self.type == "val":
            with open(self.validation_fn()) as f:
                self.raw_val_set = json.load(f)
        self._load_client = lru_cache(maxsize=512)(self._read_client)

    # -- metadata ------------------------------------------------------------------
    def _load_meta(self, train):
        with open(self.stats_fn()) as f:
            stats = json.load(f)
        self.dialogs_per_client = list(stats["dialogs_per_client"])
        self.train_utterances_per_dialog = list(stats["train_utterances_per_dialog"])
        self.val_utterances_per_dialog = list(stats["val_utterances_per_dialog"])

    def _build_index(self):
        tu = np.asarray(self.train_utterances_per_dialog, dtype=np.int64)
        self._dialog_ends = np.cumsum(tu)
        self._dialog_starts = self._dialog_ends - tu
        dpc = np.asarray(self.dialogs_per_client, dtype=np.int64)
        self._client_dialog_ends = np.cumsum(dpc)
        self._client_dialog_starts = self._client_dialog_ends - dpc
        vu = np.asarray(self.val_utterances_per_dialog, dtype=np.int64)
        self._val_ends = np.cumsum(vu)
        self._val_starts = self._val_ends - vu
        self._dpc = None
        self._client_ends = None

    @property
    def num_clients(self):
        if self.do_iid and self._num_clients is not None:
            return self._num_clients
        return len(self.dialogs_per_client)

    def _compute_data_per_client(self):
        if self.do_iid:
            total, nc = len(self), self.num_clients
            per = np.full(nc, total // nc, dtype=np.int64)
            extra = total % nc
            if extra:
                per[nc - extra:] += 1
            return per
        tu = np.asarray(self.train_utterances_per_dialog, dtype=np.int64)
        csum = np.concatenate([[0], np.cumsum(tu)])
        return csum[self._client_dialog_ends] - csum[self._client_dialog_starts]

    def __len__(self):
        per = self.train_utterances_per_dialog if self.type == "train" else self.val_utterances_per_dialog
        return int(sum(per))

    # -- items ---------------------------------------------------------------------
    def _read_client(self, client_id):
        with open(self.client_fn(client_id)) as f:
            return json.load(f)

    def __getitem__(self, idx):
        if self.type == "val":
            dialog_id = int(np.searchsorted(self._val_ends, idx, side="right"))
            dialog = self.raw_val_set[dialog_id]
            utt = dialog["utterances"][int(idx - self._val_starts[dialog_id])]
            return (-1,) + self.utterance_to_input(dialog["personality"], utt)
        src = int(self.iid_shuffle[idx]) if self.do_iid else int(idx)
        dialog_id = int(np.searchsorted(self._dialog_ends, src, side="right"))
        within_dialog = src - int(self._dialog_starts[dialog_id])
        client_id = int(np.searchsorted(self._client_dialog_ends, dialog_id, side="right"))
        dialog = self._load_client(client_id)[dialog_id - int(self._client_dialog_starts[client_id])]
        utt = dialog["utterances"][within_dialog]
        personality = list(dialog["personality"])
        item = None
        for _ in range(max(1, self.personality_permutations)):
            random.shuffle(personality)
            one = self.utterance_to_input(personality, utt)
            item = one if item is None else _merge_permutation(item, one)
        if self.do_iid:
            client_id = self._client_of_flat_index(idx)
        return (client_id,) + item

    def utterance_to_input(self, personality, utterance):
        cands = utterance["candidates"]
        n = len(cands)
        if self.num_candidates > 0 and self.type == "train":
            n = min(self.num_candidates, n)
        history = utterance["history"][-(2 * self.max_history + 1):]
        return raw_to_input(self.tokenizer, personality, history, cands[-n:])

    # -- files ---------------------------------------------------------------------
    RAW_NAME = "personachat_self_original.json"

    def prepare_datasets(self, download=True):
        """First use of a directory: split the original corpus file, if it has been placed there, into the
        per-client files (the reference downloads it first, fed_persona.py:86-126; there is no network here)."""
        raw = os.path.join(self.dataset_dir, self.RAW_NAME)
        if os.path.exists(raw):
            self.split_dataset(raw)
            return
        raise RuntimeError(
            "PersonaChat is not present in %s and cannot be downloaded (no network): put %s there, or write a "
            "synthetic corpus with data_utils.synthetic.write_synthetic_persona()" % (self.dataset_dir, self.RAW_NAME))

    def split_dataset(self, dataset_path):
        """Partition the original ``personachat_self_original.json`` by personality and write
        the per-client files (reference: fed_persona.py:128-170)."""
        with open(dataset_path) as f:
            raw = json.load(f)
        clients = defaultdict(list)
        for dialog in raw["train"]:
            clients[tuple(dialog["personality"])].append(dialog)
        dialogs_per_client, train_upd = [], []
        os.makedirs(self.dataset_dir, exist_ok=True)
        for cid, dialogs in enumerate(clients.values()):
            dialogs_per_client.append(len(dialogs))
            train_upd.extend(len(d["utterances"]) for d in dialogs)
            with open(self.client_fn(cid), "w") as f:
                json.dump(dialogs, f)
        with open(self.validation_fn(), "w") as f:
            json.dump(raw["valid"], f)
        with open(self.stats_fn(), "w") as f:
            json.dump({"dialogs_per_client": dialogs_per_client,
                       "train_utterances_per_dialog": train_upd,
                       "val_utterances_per_dialog": [len(d["utterances"]) for d in raw["valid"]]}, f)

    def client_fn(self, client_id):
        return os.path.join(self.dataset_dir, "client%d.json" % client_id)

    def validation_fn(self):
        return os.path.join(self.dataset_dir, "validation.json")


def _merge_permutation(acc, extra):
    """Concatenate the candidate lists of two renderings of the same utterance."""
    ids, mc_tok, lm, mc_label, types = acc
    ids2, mc_tok2, lm2, mc_label2, types2 = extra
    return (ids + ids2, torch.cat([mc_tok, mc_tok2]), lm + lm2, mc_label, types + types2)


def personachat_collate_fn(records):
    """List of ``(client_id,) + MODEL_INPUTS`` tuples -> tuple of padded tensors."""
    n_cand = len(records[0][1])
    batch = []
    for col, name in enumerate(["client_id"] + MODEL_INPUTS):
        if name in PADDED_INPUTS:
            seqs = [s for rec in records for s in rec[col]]
            padded = pad_sequence(seqs, batch_first=True, padding_value=-1 if name == "lm_labels" else 0)
            batch.append(padded.view(len(records), n_cand, -1))
        else:
            batch.append(torch.stack([torch.as_tensor(rec[col]) for rec in records]))
    return tuple(batch)
