"""Synthetic datasets written in the reference's own on-disk formats.

There is no network on the build or GPU boxes, so CIFAR / FEMNIST / PersonaChat
cannot be downloaded.  These writers produce *learnable* data of the same shape
(class prototype + noise images; templated persona dialogs), so accuracy-vs-round
curves are meaningful, in exactly the files ``FedCIFAR10/100``, ``FedEMNIST`` and
``FedPERSONA`` (ours and the reference's) read -- which is how the reference
baseline is benchmarked on the same data without editing it.
"""
from __future__ import annotations

import json
import os

import numpy as np
import torch

__all__ = ["write_synthetic_cifar", "write_synthetic_emnist", "write_synthetic_persona", "write_synthetic_imagenet",
           "synthetic_image_batch"]


def _prototypes(rng, num_classes, shape):
    low = rng.uniform(0, 255, size=(num_classes,) + tuple(s // 4 if i < 2 else s for i, s in enumerate(shape)))
    reps = [1, 4, 4] + ([1] if len(shape) == 3 else [])
    return np.kron(low, np.ones(reps[1:]))[:, :shape[0], :shape[1]]


def write_synthetic_cifar(dataset_dir, num_classes=10, images_per_class=500, num_val=1000, seed=0,
                          noise=48.0):
    """``client{c}.npy`` + ``test.npz`` + ``stats.json`` (fed_cifar.py formats)."""
    if os.path.exists(os.path.join(dataset_dir, "stats.json")):
        return dataset_dir
    os.makedirs(dataset_dir, exist_ok=True)
    rng = np.random.RandomState(seed)
    protos = _prototypes(rng, num_classes, (32, 32, 3))

    def draw(c, n):
        x = protos[c][None] + rng.normal(0, noise, size=(n, 32, 32, 3))
        return np.clip(x, 0, 255).astype(np.uint8)

    for c in range(num_classes):
        np.save(os.path.join(dataset_dir, "client%d.npy" % c), draw(c, images_per_class))
    ty = rng.randint(0, num_classes, size=num_val)
    tx = np.stack([draw(int(c), 1)[0] for c in ty])
    np.savez(os.path.join(dataset_dir, "test.npz"), test_images=tx, test_targets=ty)
    with open(os.path.join(dataset_dir, "stats.json"), "w") as f:
        json.dump({"images_per_client": [images_per_class] * num_classes, "num_val_images": int(num_val)}, f)
    return dataset_dir


def write_synthetic_emnist(dataset_dir, num_clients=3500, images_per_client=20, num_classes=62,
                           num_val=2000, seed=0, noise=40.0):
    """``train/client{i}.pt`` + ``test/test.pt`` + ``stats.json`` (fed_emnist.py formats);
    each writer (client) sees a skewed subset of classes, like FEMNIST."""
    if os.path.exists(os.path.join(dataset_dir, "stats.json")):
        return dataset_dir
    os.makedirs(os.path.join(dataset_dir, "train"), exist_ok=True)
    os.makedirs(os.path.join(dataset_dir, "test"), exist_ok=True)
    rng = np.random.RandomState(seed)
    protos = _prototypes(rng, num_classes, (28, 28))

    def draw(labels):
        x = protos[labels] + rng.normal(0, noise, size=(len(labels), 28, 28))
        return torch.from_numpy(np.clip(x, 0, 255).astype(np.uint8))

    counts = []
    for cid in range(num_clients):
        n = int(images_per_client)
        fav = rng.choice(num_classes, size=min(8, num_classes), replace=False)
        labels = rng.choice(fav, size=n)
        torch.save({"x": draw(labels), "y": torch.from_numpy(labels.astype(np.int64))},
                   os.path.join(dataset_dir, "train", "client%d.pt" % cid))
        counts.append(n)
    labels = rng.randint(0, num_classes, size=num_val)
    torch.save({"x": draw(labels), "y": torch.from_numpy(labels.astype(np.int64))},
               os.path.join(dataset_dir, "test", "test.pt"))
    with open(os.path.join(dataset_dir, "stats.json"), "w") as f:
        json.dump({"images_per_client": counts, "num_val_images": int(num_val)}, f)
    return dataset_dir


_TOPICS = ["dogs", "cats", "pizza", "hiking", "chess", "jazz", "soccer", "painting", "coffee", "sailing",
           "gardening", "robots", "poetry", "cycling", "baking", "astronomy"]
_VERBS = ["love", "enjoy", "dislike", "study", "collect", "teach", "watch", "dream about"]


def write_synthetic_persona(dataset_dir, num_clients=64, dialogs_per_client=2, utterances_per_dialog=4,
                            num_candidates=4, num_val_dialogs=8, seed=0, words_per_turn=12):
    """``client{i}.json`` + ``validation.json`` + ``stats.json`` (fed_persona.py formats).
    The gold reply (last candidate) mentions the persona's topic; distractors do not,
    so both the LM and the multiple-choice objective are learnable."""
    if os.path.exists(os.path.join(dataset_dir, "stats.json")):
        return dataset_dir
    os.makedirs(dataset_dir, exist_ok=True)
    rng = np.random.RandomState(seed)

    def sentence(topic, n):
        filler = ["really", "often", "today", "with", "friends", "every", "weekend", "because", "it", "is", "fun"]
        words = ["i", _VERBS[rng.randint(len(_VERBS))], topic] + [filler[rng.randint(len(filler))] for _ in range(max(0, n - 3))]
        return " ".join(words)

    def dialog(topic):
        persona = [sentence(topic, 6) for _ in range(4)]
        history, utts = [], []
        for _ in range(utterances_per_dialog):
            history = history + [sentence(_TOPICS[rng.randint(len(_TOPICS))], words_per_turn)]
            others = [t for t in _TOPICS if t != topic]
            cands = [sentence(others[rng.randint(len(others))], words_per_turn) for _ in range(num_candidates - 1)]
            gold = sentence(topic, words_per_turn)
            utts.append({"history": list(history), "candidates": cands + [gold]})
            history = history + [gold]
        return {"personality": persona, "utterances": utts}

    train_upd = []
    for cid in range(num_clients):
        topic = _TOPICS[cid % len(_TOPICS)]
        dialogs = [dialog(topic) for _ in range(dialogs_per_client)]
        train_upd.extend(len(d["utterances"]) for d in dialogs)
        with open(os.path.join(dataset_dir, "client%d.json" % cid), "w") as f:
            json.dump(dialogs, f)
    val = [dialog(_TOPICS[i % len(_TOPICS)]) for i in range(num_val_dialogs)]
    with open(os.path.join(dataset_dir, "validation.json"), "w") as f:
        json.dump(val, f)
    with open(os.path.join(dataset_dir, "stats.json"), "w") as f:
        json.dump({"dialogs_per_client": [dialogs_per_client] * num_clients,
                   "train_utterances_per_dialog": train_upd,
                   "val_utterances_per_dialog": [len(d["utterances"]) for d in val]}, f)
    return dataset_dir


def write_synthetic_imagenet(dataset_dir, num_classes=4, images_per_class=3, val_per_class=2, size=(48, 40), seed=0):
    """An ImageNet-shaped folder tree that ``torchvision.datasets.ImageNet`` (and therefore ``FedImageNet``, ours and
    the reference's: data_utils/fed_imagenet.py:11-75) opens: ``meta.bin`` plus ``train/<wnid>/*.JPEG`` and
    ``val/<wnid>/*.JPEG`` -- class-prototype + noise images, one client per wnid."""
    from PIL import Image
    rng = np.random.RandomState(seed)
    wnids = ["n%08d" % (1000 + 7 * c) for c in range(num_classes)]
    protos = rng.randint(0, 256, size=(num_classes, size[1], size[0], 3)).astype(np.float32)
    for split, per in (("train", images_per_class), ("val", val_per_class)):
        for c, wnid in enumerate(wnids):
            d = os.path.join(dataset_dir, split, wnid)
            os.makedirs(d, exist_ok=True)
            for i in range(per):
                img = np.clip(protos[c] + rng.normal(0, 24.0, protos[c].shape), 0, 255).astype(np.uint8)
                Image.fromarray(img).save(os.path.join(d, "%s_%d.JPEG" % (wnid, i)), quality=90)
    wnid_to_classes = {w: ("synthetic class %d" % c,) for c, w in enumerate(wnids)}
    val_wnids = [w for w in wnids for _ in range(val_per_class)]
    torch.save((wnid_to_classes, val_wnids), os.path.join(dataset_dir, "meta.bin"))
    return wnids


def synthetic_image_batch(num_clients, per_client, num_classes=10, shape=(3, 32, 32), seed=0,
                          first_client=0, dtype=torch.float32, pin=False, noise=0.5):
    """One round's batch ``(client_ids, images, labels)`` generated directly as tensors
    (normalised class-prototype + noise images) -- used by benchmarks and smoke tests
    where the on-disk pipeline is not the object of measurement."""
    g = torch.Generator().manual_seed(seed)
    protos = torch.randn((num_classes,) + tuple(shape), generator=torch.Generator().manual_seed(1234))
    n = num_clients * per_client
    labels = torch.randint(0, num_classes, (n,), generator=g)
    images = (protos[labels] + noise * torch.randn((n,) + tuple(shape), generator=g)).to(dtype)
    client_ids = (torch.arange(num_clients).repeat_interleave(per_client) + first_client).long()
    if pin and torch.cuda.is_available():
        images, labels, client_ids = images.pin_memory(), labels.pin_memory(), client_ids.pin_memory()
    return client_ids, images, labels
