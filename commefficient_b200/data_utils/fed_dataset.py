"""Base class of the federated datasets: a flat index space over all training
samples, a natural client partition read from ``stats.json``, and either an iid
re-partition or a split of each natural client into sub-clients.

Behavioural parity with /root/reference/CommEfficient/data_utils/fed_dataset.py:9-98
(``__getitem__`` -> ``(client_id, x, y)``, validation items carry client id -1,
``stats.json`` keys ``images_per_client`` / ``num_val_images``).  Index arithmetic
is precomputed once (the reference recomputes three cumsums per sample).
"""
from __future__ import annotations

import json
import os

import numpy as np
import torch

__all__ = ["FedDataset"]


class FedDataset(torch.utils.data.Dataset):
    def __init__(self, dataset_dir, dataset_name, transform=None, do_iid=False, num_clients=None,
                 train=True, download=False):
        self.dataset_dir = dataset_dir
        self.dataset_name = dataset_name
        self.transform = transform
        self.do_iid = bool(do_iid)
        self._num_clients = num_clients
        self.type = "train" if train else "val"
        if not self.do_iid and num_clients == 1:
            raise ValueError("a non-iid partition needs more than one client")
        if not os.path.exists(self.stats_fn()):
            self.prepare_datasets(download=download)
        self._load_meta(train)
        self._build_index()
        if self.do_iid:
            self.iid_shuffle = np.random.permutation(len(self))

    # -- partition bookkeeping -----------------------------------------------------
    def _load_meta(self, train):
        with open(self.stats_fn()) as f:
            stats = json.load(f)
        self.images_per_client = np.asarray(stats["images_per_client"], dtype=np.int64)
        self.num_val_images = int(stats["num_val_images"])

    def _build_index(self):
        self._natural_ends = np.cumsum(self.images_per_client)
        self._natural_starts = self._natural_ends - self.images_per_client
        self._dpc = None
        self._client_ends = None

    @property
    def num_clients(self):
        return self._num_clients if self._num_clients is not None else len(self.images_per_client)

    @property
    def data_per_client(self):
        if self._dpc is None:
            self._dpc = self._compute_data_per_client()
        return self._dpc

    def _compute_data_per_client(self):
        if self.do_iid:
            total, nc = len(self), self.num_clients
            per = np.full(nc, total // nc, dtype=np.int64)
            extra = total % nc
            if extra:
                per[nc - extra:] += 1
            return per
        if self._num_clients is None:
            return self.images_per_client.copy()
        # split each natural client (= class) into equally many sub-clients; the
        # remainder goes to the last sub-client of that class
        per_class = self._num_clients // len(self.images_per_client)
        if per_class < 1:
            raise ValueError("num_clients must be at least the number of natural clients")
        out = []
        for n in self.images_per_client:
            share = [n // per_class] * per_class
            share[-1] += n % per_class
            out.extend(share)
        return np.asarray(out, dtype=np.int64)

    def _client_of_flat_index(self, idx):
        if self._client_ends is None:
            self._client_ends = np.cumsum(self.data_per_client)
        return int(np.searchsorted(self._client_ends, idx, side="right"))

    # -- Dataset protocol ----------------------------------------------------------
    def __len__(self):
        return int(self.images_per_client.sum()) if self.type == "train" else self.num_val_images

    def __getitem__(self, idx):
        if self.type == "train":
            src = int(self.iid_shuffle[idx]) if self.do_iid else int(idx)
            natural = int(np.searchsorted(self._natural_ends, src, side="right"))
            image, target = self._get_train_item(natural, src - int(self._natural_starts[natural]))
            client_id = self._client_of_flat_index(idx)
        else:
            image, target = self._get_val_item(idx)
            client_id = -1
        if self.transform is not None:
            image = self.transform(image)
        return client_id, image, target

    # -- to be provided by subclasses ----------------------------------------------
    def prepare_datasets(self, download=False):
        raise NotImplementedError

    def _get_train_item(self, client_id, idx_within_client):
        raise NotImplementedError

    def _get_val_item(self, idx):
        raise NotImplementedError

    def stats_fn(self):
        return os.path.join(self.dataset_dir, "stats.json")
