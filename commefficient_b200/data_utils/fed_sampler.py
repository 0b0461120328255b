"""Round sampler: every round draws ``num_workers`` clients that still have data
and takes up to ``local_batch_size`` samples from each (all remaining samples when
``local_batch_size == -1``); an epoch ends when every client is exhausted.

Semantics of /root/reference/CommEfficient/data_utils/fed_sampler.py:5-71.  Additions used by the SPMD
runtime:

  * an explicit ``seed``: every rank constructs the identical schedule without communication;
  * ``rank`` / ``world_size``: a rank's loader only LOADS the samples of the clients that rank will simulate
    (the reference has one loader in one process; with one process per GPU an unsharded loader would decode
    and augment the whole round N times).  The schedule itself -- which clients, how many samples each --
    is still computed in full on every rank (index arithmetic only) and handed to the aggregator through
    ``rounds``, a FIFO with one ``RoundMeta`` per yielded batch, so no collective is needed to learn the
    other ranks' client ids and sample counts.
"""
from __future__ import annotations

from collections import deque
from typing import List, NamedTuple

import numpy as np

__all__ = ["FedSampler", "RoundMeta", "assign_clients", "ShardedEvalSampler", "EvalMeta"]


def assign_clients(unique_sorted: np.ndarray, world_size: int, by_owner: bool) -> List[np.ndarray]:
    """Clients (ascending ids) -> ranks.  ``by_owner``: a client always runs on the rank that owns its
    per-client state rows (id mod world size); otherwise contiguous balanced chunks.  Every client is
    processed (the reference drops W mod P clients, quirk Q1)."""
    unique_sorted = np.asarray(unique_sorted, dtype=np.int64)
    if by_owner:
        return [unique_sorted[unique_sorted % world_size == r] for r in range(world_size)]
    return [np.asarray(c, dtype=np.int64) for c in np.array_split(unique_sorted, world_size)]


class RoundMeta(NamedTuple):
    ids: np.ndarray        # the round's client ids, ascending
    counts: np.ndarray     # samples taken from each of them
    sharded: bool          # the yielded batch holds only this rank's clients (else: the whole round)


class FedSampler:
    def __init__(self, dataset, num_workers, local_batch_size, shuffle_clients=True, seed=None, rank=0,
                 world_size=1, by_owner=False):
        self.dataset = dataset
        self.num_workers = int(num_workers)
        self.local_batch_size = int(local_batch_size)
        self.shuffle_clients = shuffle_clients
        self.seed = seed
        self.rank, self.world_size, self.by_owner = int(rank), int(world_size), bool(by_owner)
        if self.world_size > 1 and seed is None:
            raise ValueError("a rank-sharded FedSampler needs a seed (all ranks must draw the same schedule)")
        self.rounds: "deque[RoundMeta]" = deque()
        self._epoch = 0

    def _rng(self):
        if self.seed is None:
            return np.random  # global numpy RNG, like the reference
        rng = np.random.RandomState(self.seed + 7919 * self._epoch)
        return rng

    def __iter__(self):
        rng = self._rng()
        self._epoch += 1
        self.rounds.clear()          # batches prefetched by an abandoned iterator are gone with it
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
            if self.world_size > 1:
                by_id = np.argsort(chosen, kind="stable")
                ids, counts = chosen[by_id], take[by_id]
                parts = assign_clients(ids, self.world_size, self.by_owner)
                sharded = all(len(p) for p in parts)     # a rank without clients still needs a non-empty batch
                self.rounds.append(RoundMeta(ids, counts, sharded))
                mine = parts[self.rank] if sharded else ids
                sel = np.searchsorted(ids, mine)
                batch = np.concatenate([order[starts[c] + used[c]: starts[c] + used[c] + t]
                                        for c, t in zip(ids[sel], counts[sel])])
            else:
                batch = np.concatenate([order[starts[c] + used[c]: starts[c] + used[c] + t]
                                        for c, t in zip(chosen, take)])
            used[chosen] += take
            yield batch

    def __len__(self):
        # number of samples, not rounds -- kept as in the reference (:70-71)
        return len(self.dataset)


class EvalMeta(NamedTuple):
    total: int             # samples of the whole validation batch
    sharded: bool          # the yielded batch holds only this rank's shards, in order (else: the whole batch)


class ShardedEvalSampler:
    """Batch sampler for validation: sequential batches of ``batch_size`` samples (what
    ``DataLoader(batch_size=valid_batch_size * num_workers, shuffle=False)`` yields in the reference's drivers),
    of which a rank only loads the ``shard_size``-sample shards it will evaluate -- the same contiguous split of the
    shard list over the ranks that ``FedModel._call_val`` uses.  ``rounds`` carries the size of every whole batch.
    With one rank this is exactly the plain sequential loader."""

    def __init__(self, num_items: int, batch_size: int, shard_size: int, rank: int = 0, world_size: int = 1):
        self.num_items, self.batch_size, self.shard_size = int(num_items), int(batch_size), int(shard_size)
        self.rank, self.world_size = int(rank), int(world_size)
        self.rounds: "deque[EvalMeta]" = deque()

    @staticmethod
    def shard_bounds(total: int, shard_size: int):
        return [(lo, min(total, lo + shard_size)) for lo in range(0, total, shard_size)]

    def __iter__(self):
        self.rounds.clear()
        for start in range(0, self.num_items, self.batch_size):
            total = min(self.batch_size, self.num_items - start)
            if self.world_size == 1:
                yield list(range(start, start + total))
                continue
            bounds = self.shard_bounds(total, self.shard_size)
            parts = np.array_split(np.arange(len(bounds)), self.world_size)
            sharded = all(len(p) for p in parts)
            self.rounds.append(EvalMeta(total, sharded))
            if not sharded:
                yield list(range(start, start + total))
                continue
            idx = []
            for si in parts[self.rank]:
                lo, hi = bounds[int(si)]
                idx.extend(range(start + lo, start + hi))
            yield idx

    def __len__(self):
        return -(-self.num_items // self.batch_size)
