"""Federated image datasets: CIFAR-10/100 (one natural client per class), FEMNIST
(one per writer) and ImageNet (one per wnid).

On-disk formats are the reference's, so prepared directories are interchangeable:
  CIFAR   ``client{i}.npy`` (uint8 N x 32 x 32 x 3), ``test.npz`` {test_images, test_targets},
          ``stats.json``            (/root/reference/CommEfficient/data_utils/fed_cifar.py:13-100)
  FEMNIST ``train/client{i}.pt`` {"x": (n,28,28), "y": (n,)}, ``test/test.pt``, ``stats.json``
          (fed_emnist.py:36-138; LEAF json is read with the stdlib, not orjson)
  ImageNet torchvision ``ImageNet`` folder layout + ``stats.json``  (fed_imagenet.py:12-76)
When the raw data cannot be downloaded (no network here) use
``data_utils.synthetic`` to write learnable data of the same shape and format.
"""
from __future__ import annotations

import json
import os
from collections import OrderedDict

import numpy as np
import torch

from .fed_dataset import FedDataset

__all__ = ["FedCIFAR10", "FedCIFAR100", "FedEMNIST", "FedImageNet"]


def _pil(array):
    from PIL import Image
    return Image.fromarray(array)


class FedCIFAR10(FedDataset):
    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        if self.type == "train":
            self.client_datasets = [np.load(self.client_fn(i)) for i in range(len(self.images_per_client))]
        else:
            with np.load(self.test_fn()) as z:
                self.test_images = z["test_images"]
                self.test_targets = z["test_targets"]

    def prepare_datasets(self, download=True):
        import torchvision
        os.makedirs(self.dataset_dir, exist_ok=True)
        cls = getattr(torchvision.datasets, self.dataset_name)
        train = cls(self.dataset_dir, train=True, download=download)
        test = cls(self.dataset_dir, train=False, download=download)
        labels = np.asarray(train.targets)
        counts = []
        for c in range(len(train.classes)):
            rows = np.flatnonzero(labels == c)
            self._write_once(self.client_fn(c), lambda fn, r=rows: np.save(fn, train.data[r]))
            counts.append(int(rows.size))
        self._write_once(self.test_fn(), lambda fn: np.savez(
            fn, test_images=test.data, test_targets=np.asarray(test.targets)))
        self._write_once(self.stats_fn(), lambda fn: json.dump(
            {"images_per_client": counts, "num_val_images": len(test.targets)}, open(fn, "w")))

    @staticmethod
    def _write_once(fn, writer):
        if os.path.exists(fn):
            raise RuntimeError("refusing to overwrite %s" % fn)
        writer(fn)

    def _get_train_item(self, client_id, idx_within_client):
        # natural client == class label
        return _pil(self.client_datasets[client_id][idx_within_client]), client_id

    def _get_val_item(self, idx):
        return _pil(self.test_images[idx]), int(self.test_targets[idx])

    def client_fn(self, client_id):
        return os.path.join(self.dataset_dir, "client%d.npy" % client_id)

    def test_fn(self):
        return os.path.join(self.dataset_dir, "test.npz")

    # -- raw-array access for the device-side fast path (no PIL) -------------------
    def raw_train_arrays(self):
        """(images uint8 [N,32,32,3], labels int64 [N]) in flat-index order."""
        images = np.concatenate(self.client_datasets, axis=0)
        labels = np.concatenate([np.full(len(a), c, dtype=np.int64) for c, a in enumerate(self.client_datasets)])
        return images, labels


class FedCIFAR100(FedCIFAR10):
    pass


def _read_leaf_json_dir(data_dir):
    users = OrderedDict()
    for name in sorted(os.listdir(data_dir)):
        if name.endswith(".json"):
            with open(os.path.join(data_dir, name)) as f:
                users.update(json.load(f)["user_data"])
    return users


class FedEMNIST(FedDataset):
    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        if self.type == "train":
            # one concatenated tensor + offsets instead of one tensor per client
            xs, ys, offsets = [], [], [0]
            for cid in range(len(self.images_per_client)):
                blob = torch.load(self.client_fn(cid))
                xs.append(blob["x"])
                ys.append(blob["y"])
                offsets.append(offsets[-1] + blob["y"].numel())
            self.client_images = torch.cat(xs, dim=0)
            self.client_targets = torch.cat(ys, dim=0)
            self.client_offsets = torch.tensor(offsets)
        else:
            blob = torch.load(self.test_fn())
            self.test_images, self.test_targets = blob["x"], blob["y"]

    def prepare_datasets(self, download=False):
        for fn in (self.stats_fn(), self.test_fn()):
            if os.path.exists(fn):
                raise RuntimeError("refusing to overwrite %s" % fn)
        counts = []
        train_users = _read_leaf_json_dir(os.path.join(self.dataset_dir, "train"))
        for cid, blob in enumerate(train_users.values()):
            x = torch.tensor(blob["x"]).view(-1, 28, 28)
            y = torch.tensor(blob["y"])
            counts.append(y.numel())
            if not os.path.exists(self.client_fn(cid)):
                torch.save({"x": x, "y": y}, self.client_fn(cid))
        test_users = _read_leaf_json_dir(os.path.join(self.dataset_dir, "test"))
        tx = torch.cat([torch.tensor(b["x"]).view(-1, 28, 28) for b in test_users.values()])
        ty = torch.cat([torch.tensor(b["y"]) for b in test_users.values()])
        torch.save({"x": tx, "y": ty}, self.test_fn())
        with open(self.stats_fn(), "w") as f:
            json.dump({"images_per_client": counts, "num_val_images": int(ty.numel())}, f)

    def _get_train_item(self, client_id, idx_within_client):
        pos = int(self.client_offsets[client_id]) + int(idx_within_client)
        return _pil(self.client_images[pos].numpy()), int(self.client_targets[pos])

    def _get_val_item(self, idx):
        return _pil(self.test_images[idx].numpy()), int(self.test_targets[idx])

    def test_fn(self):
        return os.path.join(self.dataset_dir, "test", "test.pt")

    def client_fn(self, client_id):
        return os.path.join(self.dataset_dir, "train", "client%d.pt" % client_id)

    def raw_train_arrays(self):
        return self.client_images.numpy(), self.client_targets.numpy().astype(np.int64)


class FedImageNet(FedDataset):
    """One client per wnid over an already-extracted ImageNet folder."""

    def __init__(self, *args, **kwargs):
        if kwargs.get("download"):
            raise RuntimeError("ImageNet cannot be downloaded; point --dataset_dir at an extracted copy")
        super().__init__(*args, **kwargs)
        from torchvision.datasets import ImageNet
        self.vanilla_train = ImageNet(self.dataset_dir, split="train")
        self.vanilla_val = ImageNet(self.dataset_dir, split="val")

    def prepare_datasets(self, download=False):
        if download:
            raise RuntimeError("ImageNet cannot be downloaded")
        from torchvision.datasets import ImageNet
        train = ImageNet(self.dataset_dir, split="train")
        val = ImageNet(self.dataset_dir, split="val")
        labels = np.asarray([t for _, t in train.samples])
        # samples are grouped by class in folder order
        change = np.flatnonzero(np.diff(labels)) + 1
        bounds = np.concatenate([[0], change, [len(labels)]])
        counts = np.diff(bounds).tolist()
        if os.path.exists(self.stats_fn()):
            raise RuntimeError("refusing to overwrite %s" % self.stats_fn())
        with open(self.stats_fn(), "w") as f:
            json.dump({"images_per_client": counts, "num_val_images": len(val.samples)}, f)

    def _get_train_item(self, client_id, idx_within_client):
        return self.vanilla_train[int(self._natural_starts[client_id]) + int(idx_within_client)]

    def _get_val_item(self, idx):
        return self.vanilla_val[idx]
