"""Federated facade: ``FedModel`` (callable that runs one round of clients) and
``FedOptimizer`` (server update), with the calling convention of the reference
(/root/reference/CommEfficient/fed_aggregator.py:54-616):

    model = FedModel(module, compute_loss_train, args, compute_loss_val)
    opt = FedOptimizer(torch.optim.SGD(params, lr=1), args)
    sched = LambdaLR(opt, ...)
    loss, acc, download, upload = model(batch); opt.step()

Architecture: SPMD -- every rank runs this same code.  A rank
simulates its share of the round's clients (``fed_worker.ClientWorker``), the per-rank
transmit buffers are summed across ranks (peer-memory kernels / NCCL / gloo), and every
rank then applies the *same* deterministic server update to its own replica of the
flat weights, so the weights never have to be broadcast: the reference's host-shm
weight download (X1), device->host update upload (X2) and multiprocessing queues (X4/X5)
have no counterpart.  Server state (``Vvelocity`` / ``Verror``) and weights live on
the device; nothing in a round synchronises with the host except one small D2H copy
of the per-client results.
"""
from __future__ import annotations

import math
import os
import warnings
from typing import List, Optional

import numpy as np
import torch

from . import fed_worker as worker
from .data_utils.fed_sampler import assign_clients
from .ops import dense as dn
from .ops import sketch as sk
from .ops.topk import topk_sparse
from .parallel.runtime import Runtime, init_runtime
from .parallel.watchdog import RoundWatchdog, maybe_inject, parse_fault
from .utils import graphs
from .utils.logging import PhaseTracer
from .utils.flat import get_param_vec, set_param_vec  # noqa: F401  (re-exported, reference names)

DEFAULT_NUM_CLIENTS = {"EMNIST": 3500, "PERSONA": 17568}

# The FedModel most recently constructed in this process; FedOptimizer binds to it the
# way the reference couples the two through module globals (fed_aggregator.py:37-44).
_CURRENT: Optional["FedModel"] = None


def _upload_floats(args, d: int) -> int:
    return {"uncompressed": d, "true_topk": d, "local_topk": args.k,
            "sketch": args.num_rows * args.num_cols, "fedavg": d}[args.mode]


class _SmallH2D:
    """Asynchronous host->device copies of small per-round index arrays.  ``torch.as_tensor(np_array,
    device="cuda")`` copies from pageable memory, which blocks the host until everything queued on the
    stream before it -- i.e. the whole client round -- has finished; staging through a ring of pinned
    buffers (each guarded by the event of its last copy) keeps the host ahead of the GPU."""

    def __init__(self, device, slots: int = 8):
        self.device = torch.device(device)
        self.slots, self.next = slots, 0
        self.ring = {}

    def __call__(self, arr: np.ndarray) -> torch.Tensor:
        arr = np.ascontiguousarray(arr)
        if self.device.type != "cuda":
            return torch.as_tensor(arr, device=self.device)
        key = (arr.dtype.str, arr.size)
        ring = self.ring.setdefault(key, [])
        if len(ring) < self.slots:
            ring.append([torch.empty(arr.shape, dtype=torch.from_numpy(arr).dtype).pin_memory(), None])
            slot = ring[-1]
        else:
            slot = ring[self.next % self.slots]
            self.next += 1
            if slot[1] is not None:
                slot[1].synchronize()           # the copy that last used this buffer has long finished
        slot[0].numpy()[...] = arr.reshape(slot[0].shape)
        dev = slot[0].to(self.device, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        slot[1] = ev
        return dev


class _DownloadTracker:
    """Bytes a client must download = 4 x (coordinates changed since it last took part).
    Kept on the device as ``last_changed[d]`` (round of each coordinate's last update)
    plus a histogram over rounds, so a client's count is a suffix sum -- instead of the
    reference's deque of full weight copies and W full-vector diffs on the host per
    round (fed_aggregator.py:239-289, quirk Q19)."""

    def __init__(self, d: int, num_clients: int, device, sparse: bool):
        self.d, self.device, self.sparse = d, device, sparse
        self.round = 0                      # number of server updates applied so far
        # (on CUDA the histogram is sized so that growing it -- which moves it and invalidates the server
        # step's CUDA graph -- practically never happens)
        self.hist = torch.zeros(65536 if torch.device(device).type == "cuda" else 1024, dtype=torch.int64,
                                device=device)
        self.h2d = _SmallH2D(device)
        self.generation = 0                 # bumped whenever hist / last are re-allocated
        self.round_dev = None               # device copy of `round` for graph-replayed updates
        self.last = torch.zeros(d, dtype=torch.int64, device=device) if sparse else None
        self.hist[0] = d                    # bucket 0 = never changed; never inside a suffix sum
        self.seen = np.zeros(num_clients, dtype=np.int64)   # server round each client last saw

    def _grow(self, need: int):
        if need >= self.hist.numel():
            new = torch.zeros(max(need + 1, 2 * self.hist.numel()), dtype=torch.int64, device=self.device)
            new[: self.hist.numel()] = self.hist
            self.hist = new
            self.generation += 1

    def download_counts(self, clients: np.ndarray) -> torch.Tensor:
        """Device float32 tensor: changed-coordinate count per participating client."""
        start_np = np.minimum(self.seen[clients] + 1, self.hist.numel() - 1)
        self.seen[clients] = self.round
        start = self.h2d(start_np.astype(np.int64))
        if self.hist.is_cuda:
            from .ops import native
            out = torch.empty(len(clients), dtype=torch.float32, device=self.device)
            native.call("ceb_tracker_counts", native.ptr(self.hist), self.hist.numel(), native.ptr(start),
                        len(clients), native.ptr(out), native.stream_ptr(self.device))
            return out
        suffix = torch.flip(torch.cumsum(torch.flip(self.hist, [0]), 0), [0])
        return suffix[start].to(torch.float32)

    def record_sparse(self, idx: torch.Tensor):
        self.round += 1
        self._grow(self.round)
        if idx.is_cuda and idx.dtype == torch.int32 and idx.is_contiguous():
            from .ops import native
            native.call("ceb_tracker_record", native.ptr(idx), idx.numel(), native.ptr(self.last),
                        native.ptr(self.hist), self.round, native.stream_ptr(self.device))
            return
        li = idx.long()
        old = self.last[li]
        self.hist.index_add_(0, old, torch.full_like(old, -1))
        self.hist[self.round] += li.numel()
        self.last[li] = self.round

    # -- CUDA-graph friendly variant: the round number lives on the device -----------------
    def sync_round_dev(self):
        if self.round_dev is None:
            self.round_dev = torch.zeros(1, dtype=torch.int64, device=self.device)
        self.round_dev.fill_(self.round)

    def host_advance(self):
        """Host half of ``record_sparse_graph``: advance the round and make room in the histogram."""
        self.round += 1
        self._grow(self.round)

    def record_sparse_graph(self, idx: torch.Tensor):
        """Device half (capturable): ``round_dev += 1`` then move ``idx`` into that round's bucket."""
        from .ops import native
        self.round_dev.add_(1)
        native.call("ceb_tracker_record_dev", native.ptr(idx), idx.numel(), native.ptr(self.last),
                    native.ptr(self.hist), native.ptr(self.round_dev), native.stream_ptr(self.device))

    def record_dense(self):
        self.round += 1
        self._grow(self.round)
        self.hist.zero_()
        self.hist[self.round] = self.d


class _DeferredRound:
    """One round's packed per-client results on their way to the host (``--defer_results``): an async
    copy into a pinned buffer plus the event that completes it.  ``get()`` blocks on the event the
    first time anything looks at a value."""

    def __init__(self, host, event, watchdog=None, peer=None, peer_flag=None):
        self.host, self.event, self.watchdog = host, event, watchdog
        self.peer, self.peer_flag = peer, peer_flag
        self.array = None

    def get(self) -> np.ndarray:
        if self.array is None:
            if self.event is not None:
                self.event.synchronize()
            if self.peer_flag is not None and float(self.peer_flag[0]) != 0.0:
                self.peer.check()          # a device-side barrier of an earlier collective gave up: raise it here
            self.array = self.host.numpy().copy()
            if self.watchdog is not None:
                self.watchdog.idle()
        return self.array


class DeferredArray:
    """numpy-like view of one result column of a ``_DeferredRound``: converts itself on first use
    (``np.mean(x)``, ``np.isnan(x)``, iteration, indexing, arithmetic, any ndarray attribute)."""

    def __init__(self, rnd: _DeferredRound, fn):
        self._rnd, self._fn, self._val = rnd, fn, None

    def materialize(self) -> np.ndarray:
        if self._val is None:
            self._val = self._fn(self._rnd.get())
        return self._val

    def __array__(self, dtype=None, copy=None):
        a = self.materialize()
        return a.astype(dtype) if dtype is not None else a

    def __len__(self):
        return len(self.materialize())

    def __iter__(self):
        return iter(self.materialize())

    def __getitem__(self, i):
        return self.materialize()[i]

    def __getattr__(self, name):
        return getattr(self.materialize(), name)

    def __repr__(self):
        return "DeferredArray(%r)" % (self.materialize(),)


def _binop(name):
    def op(self, other):
        other = other.materialize() if isinstance(other, DeferredArray) else other
        return getattr(self.materialize(), name)(other)
    return op


for _n in ("__add__", "__radd__", "__sub__", "__rsub__", "__mul__", "__rmul__", "__truediv__", "__rtruediv__",
           "__lt__", "__le__", "__gt__", "__ge__", "__eq__", "__ne__"):
    setattr(DeferredArray, _n, _binop(_n))
DeferredArray.__hash__ = object.__hash__


class DeferredTensor:
    """torch-side counterpart for the download / upload byte vectors: any torch function that
    receives one (``total += dl``) sees the materialised CPU tensor."""

    def __init__(self, rnd: _DeferredRound, fn):
        self._rnd, self._fn, self._val = rnd, fn, None

    def materialize(self) -> torch.Tensor:
        if self._val is None:
            self._val = self._fn(self._rnd.get())
        return self._val

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        def conv(a):
            if isinstance(a, DeferredTensor):
                return a.materialize()
            if isinstance(a, (list, tuple)):
                return type(a)(conv(x) for x in a)
            return a
        return func(*conv(args), **{k: conv(v) for k, v in (kwargs or {}).items()})

    def __getattr__(self, name):
        return getattr(self.materialize(), name)

    def __getitem__(self, i):
        return self.materialize()[i]

    def __array__(self, dtype=None, copy=None):
        a = self.materialize().numpy()
        return a.astype(dtype) if dtype is not None else a


class FedModel:
    def __init__(self, input_model, compute_loss, args, compute_loss_val=None, runtime: Optional[Runtime] = None):
        global _CURRENT
        self.args = args
        self.rt = runtime or init_runtime(args)
        self.device = self.rt.device
        num_clients = args.num_clients
        if num_clients is None:
            num_clients = DEFAULT_NUM_CLIENTS.get(args.dataset_name)
            if num_clients is None:
                raise ValueError("--num_clients is required for dataset %r" % (args.dataset_name,))
        self.num_clients = int(num_clients)
        self.model = input_model
        self.compute_loss_train = compute_loss
        self.compute_loss_val = compute_loss_val if compute_loss_val is not None else compute_loss
        self.worker = worker.ClientWorker(input_model, args, self.compute_loss_train, self.compute_loss_val,
                                          self.device, self.num_clients, peer=self.rt.peer)
        # With the peer-memory backend the cross-rank sum is fused into the server kernel
        # (FedOptimizer.step); otherwise it happens here, right after the clients ran.
        self.fused_reduce = (self.rt.peer is not None
                             and args.mode in ("sketch", "uncompressed", "local_topk", "fedavg", "true_topk")
                             and not (args.mode == "uncompressed" and args.do_dp and args.dp_mode == "server")
                             and not getattr(args, "p2", 0))
        self.reduce_pending = False
        self.fp = self.worker.fp
        if self.rt.world_size > 1:
            # replicas must start identical: take rank 0's initial weights
            self.rt.broadcast_(self.fp.flat, 0)
        self.training = True
        self.pending_samples = 0          # N of the round whose transmit is waiting for opt.step()
        self.fedavg_lr = 0.0              # published by FedOptimizer.step (fed_aggregator.py:441-444)
        self.rounds = 0
        self.tracker = _DownloadTracker(self.fp.numel, self.num_clients, self.device,
                                        sparse=args.mode in ("sketch", "true_topk"))
        self.last_unique_clients = np.zeros(0, dtype=np.int64)
        self.watchdog = RoundWatchdog(getattr(args, "watchdog_timeout", 0.0) or 0.0, self.rt.rank).start()
        self._fault = parse_fault()
        self._result_ring = {}
        self._copy_stream = None
        self.round_source = None          # a rank-sharded FedSampler (attach_sampler)
        self.val_source = None            # a rank-sharded ShardedEvalSampler (attach_val_sampler)
        self.tracer = PhaseTracer(getattr(args, "trace_phases", "") or "", getattr(args, "trace_every", 50),
                                  self.rt.rank, self._trace_max, self.device.type)
        _CURRENT = self

    def _trace_max(self, vals):
        if self.rt.world_size == 1:
            return vals
        return [self.rt.max_over_ranks(v) for v in vals]

    # -- nn.Module-ish surface -------------------------------------------------------
    def train(self, training=True):
        self.training = bool(training)
        return self

    def eval(self):
        return self.train(False)

    def finalize(self):
        """End of training (the reference joins its worker processes here)."""
        self.watchdog.stop()
        self.tracer.flush()
        self.rt.barrier()

    def zero_grad(self):
        warnings.warn("gradients are zeroed before every client's backward pass")
        self.fp.zero_grad()

    def parameters(self):
        return self.model.parameters()

    def state_dict(self):
        # parameters are views of the flat weights: always current
        return self.model.state_dict()

    def load_state_dict(self, state, strict=True):
        out = self.model.load_state_dict(state, strict=strict)
        self.fp.rebind()
        return out

    def save_pretrained(self, log_dir):
        if self.rt.rank == 0:
            self.model.save_pretrained(log_dir)

    def __getattr__(self, name):
        # anything else is forwarded to the wrapped module
        model = self.__dict__.get("model")
        if model is not None and hasattr(model, name):
            return getattr(model, name)
        raise AttributeError(name)

    def __call__(self, batch):
        return self._call_train(batch) if self.training else self._call_val(batch)

    # -- one training round ------------------------------------------------------------
    def _assign(self, unique: np.ndarray) -> List[np.ndarray]:
        """Clients -> ranks.  With per-client state a client always runs on the rank that
        owns its rows (client id mod world size); otherwise contiguous balanced chunks.
        Every client is processed (the reference drops W mod P clients, quirk Q1)."""
        return assign_clients(unique, self.rt.world_size, self._by_owner())

    def _by_owner(self) -> bool:
        w = self.worker
        return w.velocities is not None or w.errors is not None or w.client_weights is not None

    def attach_sampler(self, sampler) -> None:
        """Tell the model that its training batches come from a rank-sharded ``FedSampler`` (one whose loader
        only materialises this rank's clients): the round's global client ids / sample counts are then read
        from ``sampler.rounds`` instead of from the batch.  Call before the first iteration."""
        if self.rt.world_size > 1 and getattr(sampler, "seed", None) is None:
            raise ValueError("attach_sampler: a rank-sharded FedSampler needs a seed (every rank must draw the same rounds)")
        sampler.rank, sampler.world_size = self.rt.rank, self.rt.world_size
        sampler.by_owner = self._by_owner()
        self.round_source = sampler if self.rt.world_size > 1 else None

    def _to_device(self, t: torch.Tensor) -> torch.Tensor:
        """Inputs from pinned host memory are copied on a dedicated stream, so the transfer of round t+1
        starts as soon as the driver calls the model -- under round t's server step -- instead of queueing
        behind it; the compute stream waits on the copy's event."""
        if t.device == self.device:
            return t
        if self.device.type != "cuda" or not t.is_pinned():
            return t.to(self.device, non_blocking=True)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
        main = torch.cuda.current_stream(self.device)
        with torch.cuda.stream(self._copy_stream):
            dev = t.to(self.device, non_blocking=True)
        main.wait_stream(self._copy_stream)
        dev.record_stream(main)
        return dev

    def attach_val_sampler(self, sampler) -> None:
        """Validation batches come from a ``ShardedEvalSampler``: each rank loads only the shards it evaluates."""
        sampler.rank, sampler.world_size = self.rt.rank, self.rt.world_size
        self.val_source = sampler if self.rt.world_size > 1 else None

    def val_batch_size(self, batch) -> int:
        """Samples of the WHOLE validation batch ``batch`` belongs to (see ``round_shape``)."""
        src = self.val_source
        if src is not None and src.rounds:
            return int(src.rounds[0].total)
        return int(batch[0].shape[0])

    def skip_val_batch(self) -> None:
        src = self.val_source
        if src is not None and src.rounds:
            src.rounds.popleft()

    def round_shape(self, batch):
        """``(clients, samples)`` of the WHOLE round ``batch`` belongs to.  With a rank-sharded sampler the batch
        holds only this rank's share, so a driver that decides "incomplete round, skip it" (the reference's loop:
        cv_train.py:196-204) must ask here, not count the batch."""
        src = self.round_source
        if src is not None and src.rounds:
            meta = src.rounds[0]
            return len(meta.ids), int(meta.counts.sum())
        ids = batch[0]
        ids = ids if torch.is_tensor(ids) else torch.as_tensor(np.asarray(ids))
        return int(torch.unique(ids).numel()), int(ids.numel())

    def skip_round(self) -> None:
        """The driver drops the batch it just fetched without running it: drop its schedule entry too."""
        src = self.round_source
        if src is not None and src.rounds:
            src.rounds.popleft()

    def check_replicas(self) -> None:
        """Raise if the ranks' weight replicas are not bit-identical.  Every rank applies the same deterministic
        server update to its own replica and weights are never re-broadcast (module docstring), so a divergence --
        a non-deterministic tie-break, a missed collective -- would otherwise go unnoticed."""
        if not self.rt.replicas_agree(self.fp.flat):
            raise RuntimeError("weight replicas differ across ranks after %d rounds (rank %d)" % (self.rounds, self.rt.rank))

    def _call_train(self, batch):
        args = self.args
        every = int(getattr(args, "replica_check_every", 0) or 0)
        if every > 0 and self.rounds > 0 and self.rounds % every == 0:
            self.check_replicas()
        ids_t = batch[0]
        ids = ids_t.cpu().numpy() if torch.is_tensor(ids_t) else np.asarray(ids_t)
        sorted_already = bool(ids.size <= 1 or np.all(ids[1:] >= ids[:-1]))
        if sorted_already:
            # the samplers emit each client's samples together in ascending client order: group
            # boundaries are where the id changes (no sort / inverse permutation needed)
            starts = np.concatenate([[0], np.flatnonzero(ids[1:] != ids[:-1]) + 1])
            unique = ids[starts]
            counts = np.diff(np.concatenate([starts, [ids.size]]))
            order = None
        else:
            unique, inverse = np.unique(ids, return_inverse=True)
            order = np.argsort(inverse, kind="stable")
            counts = np.bincount(inverse, minlength=len(unique))
            starts = np.concatenate([[0], np.cumsum(counts)[:-1]])
        # rows of `starts` / `counts` (this batch's clients) that this rank simulates, and the round as a whole
        total_samples = int(ids.shape[0])
        meta = None
        if getattr(self, "round_source", None) is not None:
            if not self.round_source.rounds:
                raise RuntimeError("attach_sampler: the sampler has no pending round for this batch")
            meta = self.round_source.rounds.popleft()
        if meta is not None and meta.sharded:
            # the loader produced only this rank's clients; the schedule gives the rest of the round
            mine = self._assign(meta.ids)[self.rt.rank]
            if not np.array_equal(np.sort(unique), mine):
                raise RuntimeError("rank-sharded batch does not hold the clients assigned to rank %d" % self.rt.rank)
            local_unique, unique = unique, meta.ids
            total_samples = int(meta.counts.sum())
            pos = np.searchsorted(local_unique, mine)
        else:
            mine = self._assign(unique)[self.rt.rank]
            pos = np.searchsorted(unique, mine)
        cols = batch[1:]
        client_batches = worker.ClientBatches()
        tr = self.tracer
        h2d_phase = tr.phase("h2d")
        h2d_phase.__enter__()
        if len(mine):
            if sorted_already and int(pos[-1] - pos[0]) + 1 == len(pos):
                # this rank's clients are one contiguous run of samples: slice (views of the
                # caller's -- ideally pinned -- tensors), then one async H2D per column
                a, b = int(starts[pos[0]]), int(starts[pos[-1]] + counts[pos[-1]])
                dev_cols = [self._to_device(t[a:b]) for t in cols]
            else:
                if order is None:
                    order = np.arange(ids.size)
                sel = np.concatenate([order[starts[p]:starts[p] + counts[p]] for p in pos])
                sel_t = torch.from_numpy(sel)
                dev_cols = [self._to_device(t[sel_t] if t.device.type == "cpu" else t[sel_t.to(t.device)])
                            for t in cols]
            client_batches = worker.ClientBatches(mine, counts[pos], dev_cols)
        h2d_phase.__exit__(None, None, None)
        self.watchdog.beat("clients", self.rounds + 1)
        with tr.phase("clients"):
            local = self.worker.run_round(client_batches, fedavg_lr=self.fedavg_lr)
        maybe_inject(self._fault, self.rt.rank, self.rounds + 1, self.worker.transmit)
        self.watchdog.beat("reduce")
        if self.fused_reduce:
            self.reduce_pending = True
        else:
            with tr.phase("reduce"):
                self.rt.all_reduce_sum_(self.worker.transmit)
        self.pending_samples = total_samples
        self.last_unique_clients = unique
        self.rounds += 1

        # per-client results of every rank, in ascending client-id order
        res_phase = tr.phase("results")
        res_phase.__enter__()
        dl = self.tracker.download_counts(unique).to(torch.float32)
        if self.rt.world_size > 1 and self.rt.peer is not None:
            # every rank scatters its clients' rows into a zeroed [W, n_res] slab and the slabs are
            # summed with one small peer-memory kernel (no NCCL launch sequence, no padding traffic)
            slab = torch.zeros((len(unique), local.shape[1] if local.dim() == 2 else args.num_results_train),
                               dtype=torch.float32, device=self.device)
            if len(mine):
                slab[self.tracker.h2d(np.searchsorted(unique, mine).astype(np.int64))] = local.float()
            gathered = self.rt.peer.all_reduce_sum_(slab)
        elif self.rt.world_size > 1:
            sizes = [len(c) for c in self._assign(unique)]
            gathered = self.rt.all_gather_cat(local.float(), sizes)
            perm = np.argsort(np.concatenate(self._assign(unique)), kind="stable")
            gathered = gathered[torch.from_numpy(perm).to(gathered.device)]
        else:
            gathered = local.float()
        self.watchdog.beat("results")
        packed_dev = torch.cat([gathered, dl.unsqueeze(1)], dim=1)
        n_res = args.num_results_train
        n_cl = self.num_clients
        up_floats = 4.0 * _upload_floats(args, self.fp.numel)
        uniq_t = torch.from_numpy(unique)

        def _download(packed):
            out = torch.zeros(n_cl)
            out[uniq_t] = torch.from_numpy(4.0 * packed[:, -1]).float()
            return out

        def _upload(_packed):
            out = torch.zeros(n_cl)
            out[uniq_t] = up_floats
            return out

        if getattr(args, "defer_results", False) and packed_dev.is_cuda:
            # the round's one D2H is asynchronous: the caller may queue the server step (opt.step())
            # before anything waits for the values, so the GPU never idles behind the host
            host = self._result_buffer(packed_dev.shape)
            host.copy_(packed_dev, non_blocking=True)
            peer, flag = self.rt.peer, None
            if peer is not None:
                # the error word of the peer-memory barriers rides along with the round's one D2H
                flag = self._result_buffer((1,))
                flag.copy_((peer.ticket[2:3] != 0).float(), non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
            res_phase.__exit__(None, None, None)
            rnd = _DeferredRound(host, ev, self.watchdog, peer, flag)
            results = [DeferredArray(rnd, (lambda pk, i=i: pk[:, i].copy())) for i in range(n_res)]
            return results + [DeferredTensor(rnd, _download), DeferredTensor(rnd, _upload)]
        packed = packed_dev.cpu().numpy()   # the round's one D2H
        if self.rt.peer is not None:
            self.rt.peer.check()            # (the device is idle here: reading the barrier error word costs nothing)
        res_phase.__exit__(None, None, None)
        self.watchdog.idle()
        results = [packed[:, i].copy() for i in range(n_res)]
        return results + [_download(packed), _upload(packed)]

    def _result_buffer(self, shape):
        """Pinned host buffers for deferred results, rotated so that a round's values stay valid
        until two further rounds have been issued."""
        key = tuple(shape)
        ring = self._result_ring.setdefault(key, {"bufs": [], "next": 0})
        if len(ring["bufs"]) < 3:
            ring["bufs"].append(torch.empty(key, dtype=torch.float32, pin_memory=True))
            return ring["bufs"][-1]
        buf = ring["bufs"][ring["next"]]
        ring["next"] = (ring["next"] + 1) % 3
        return buf

    # -- validation ----------------------------------------------------------------------
    def _call_val(self, batch):
        args = self.args
        n = batch[0].shape[0]
        vb = args.valid_batch_size
        meta = None
        if self.val_source is not None:
            if not self.val_source.rounds:
                raise RuntimeError("attach_val_sampler: the sampler has no pending entry for this batch")
            meta = self.val_source.rounds.popleft()
            n = int(meta.total)
        shard_bounds = [(lo, min(n, lo + vb)) for lo in range(0, n, vb)]
        mine = np.array_split(np.arange(len(shard_bounds)), self.rt.world_size)[self.rt.rank]
        shards = []
        offset = 0
        for si in mine:
            lo, hi = shard_bounds[si]
            if meta is not None and meta.sharded:
                # the loader produced only this rank's shards, one after the other
                lo, hi = offset, offset + (hi - lo)
                offset = hi
            shards.append([self._to_device(t[lo:hi]) for t in batch[1:]])
        if meta is not None and meta.sharded and offset != batch[0].shape[0]:
            raise RuntimeError("rank-sharded validation batch has %d rows, expected %d" % (batch[0].shape[0], offset))
        local = self.worker.run_val(shards)
        if self.rt.world_size > 1:
            sizes = [len(c) for c in np.array_split(np.arange(len(shard_bounds)), self.rt.world_size)]
            local = self.rt.all_gather_cat(local.float(), sizes)
        out = local.float().cpu().numpy()
        return [out[:, i].copy() for i in range(args.num_results_val)]

    # -- resume sidecar ----------------------------------------------------
    def fedstate(self, opt: "FedOptimizer", scheduler=None):
        w = self.worker
        return {
            "weights": self.fp.flat.detach().cpu(),
            "Vvelocity": opt.Vvelocity.detach().cpu(), "Verror": opt.Verror.detach().cpu(),
            "rounds": self.rounds, "fedavg_lr": self.fedavg_lr,
            "tracker": {"round": self.tracker.round, "hist": self.tracker.hist.cpu(),
                        "last": None if self.tracker.last is None else self.tracker.last.cpu(),
                        "seen": self.tracker.seen.copy()},
            "client_velocities": None if w.velocities is None else w.velocities.state_dict(),
            "client_errors": None if w.errors is None else w.errors.state_dict(),
            "client_weights": None if w.client_weights is None else w.client_weights.state_dict(),
            "torch_rng": torch.random.get_rng_state(), "numpy_rng": np.random.get_state(),
            # what the flat vectors mean: a state saved under one layout / world size / sharding must not be
            # loaded under another (conv weights would be silently permuted, shards misplaced)
            "flat_layout": self.fp.layout, "numel": int(self.fp.numel), "world_size": self.rt.world_size,
            "rank": self.rt.rank, "shard": (int(opt.shard_lo), int(opt.shard_hi)), "mode": self.args.mode,
            "sched_step": None if scheduler is None else int(scheduler.last_epoch),
        }

    def _fedstate_path(self, path):
        return "%s.rank%d" % (path, self.rt.rank) if self.rt.world_size > 1 else path

    def save_fedstate(self, path, opt: "FedOptimizer", scheduler=None):
        """Resume sidecar.  EVERY rank must call this: with more than one rank each writes
        ``<path>.rank<r>`` (its Vvelocity shard and the per-client rows it owns live nowhere else)."""
        torch.save(self.fedstate(opt, scheduler), self._fedstate_path(path))
        self.rt.barrier()

    def load_fedstate(self, path, opt: "FedOptimizer", scheduler=None):
        st = torch.load(self._fedstate_path(path), map_location="cpu", weights_only=False)
        for key, have in (("flat_layout", self.fp.layout), ("numel", int(self.fp.numel)),
                          ("world_size", self.rt.world_size), ("mode", self.args.mode)):
            if key in st and st[key] != have:
                raise ValueError("resume state %s was saved with %s=%r, this run has %r"
                                 % (path, key, st[key], have))
        self.fp.flat.copy_(st["weights"].to(self.device))
        opt.Vvelocity.copy_(st["Vvelocity"].to(self.device))
        opt.Verror.copy_(st["Verror"].to(self.device))
        self.rounds, self.fedavg_lr = st["rounds"], st["fedavg_lr"]
        self.worker.round_no = int(st["rounds"])
        tr = st["tracker"]
        self.tracker.round, self.tracker.seen = tr["round"], tr["seen"]
        self.tracker.hist = tr["hist"].to(self.device)
        if tr["last"] is not None:
            self.tracker.last = tr["last"].to(self.device)
        self.tracker.generation += 1          # re-allocated: any captured server-step graph is stale
        w = self.worker
        for store, key in ((w.velocities, "client_velocities"), (w.errors, "client_errors"),
                           (w.client_weights, "client_weights")):
            if store is not None and st[key] is not None:
                store.load_state_dict(st[key])
        torch.random.set_rng_state(st["torch_rng"])
        np.random.set_state(st["numpy_rng"])
        self.resumed_sched_step = st.get("sched_step")
        if scheduler is not None:
            self.restore_scheduler(scheduler)

    def restore_scheduler(self, scheduler):
        """Put a ``LambdaLR`` back at the step the sidecar was saved at (the drivers build the
        scheduler after loading the state); without this the learning rate would restart at 0."""
        step = getattr(self, "resumed_sched_step", None)
        if step is None:
            return
        scheduler.last_epoch = int(step)
        lrs = []
        for group, base, lmbda in zip(scheduler.optimizer.param_groups, scheduler.base_lrs, scheduler.lr_lambdas):
            group["lr"] = base * lmbda(int(step))
            lrs.append(group["lr"])
        scheduler._last_lr = lrs


class FedOptimizer(torch.optim.Optimizer):
    """Server-side optimiser.  Borrows the ``param_groups`` of a wrapped ``SGD(lr=1)`` so a
    ``LambdaLR`` can drive the learning rate (the scheduled factor *is* the LR, as in the
    reference, fed_aggregator.py:383-461) and owns the server momentum / error state."""

    def __init__(self, optimizer, args, fed_model: Optional[FedModel] = None):
        self.fed_model = fed_model or _CURRENT
        if self.fed_model is None:
            raise RuntimeError("construct FedModel before FedOptimizer")
        self.args = args
        super().__init__(optimizer.param_groups, {})
        fm = self.fed_model
        device = fm.device
        shape = (args.num_rows, args.num_cols) if args.mode == "sketch" else (fm.fp.numel,)
        peer = fm.rt.peer
        numel = int(np.prod(shape))
        self.Vvelocity = torch.zeros(numel + 4, dtype=torch.float32, device=device)[:numel].view(shape)
        self.verror_buf = None
        if fm.fused_reduce and args.mode == "sketch":
            # Verror lives in symmetric memory: the fused kernel all-gathers the updated error
            # table straight into every rank's copy.  Vvelocity is only maintained on this
            # rank's shard [shard_lo, shard_hi) of the flattened table.
            self.verror_buf = peer.symm_empty(peer.padded(numel))
            self.Verror = self.verror_buf.tensor[:numel].view(shape)
        else:
            self.Verror = torch.zeros(shape, dtype=torch.float32, device=device)
        self.shard_lo, self.shard_hi = (peer.shard_range(peer.padded(numel)) if fm.fused_reduce else (0, numel))
        self._lr_vec = None
        self._lr_vec_key = None
        self._noise_calls = 0
        self.last_update = None           # (idx, val) of the last sparse update, for inspection
        self._tail = {"calls": 0, "graph": None, "gen": -1, "src": 0, "lr": None, "out": None, "native": 0}
        # multi-GPU sparse modes: the server tail is sharded -- every rank estimates / owns 1/W of the coordinates
        # and the exact top-k is found by the distributed select (parallel/symm.py:DistSelect)
        self.dsel = None
        self._est_slice = None
        if fm.fused_reduce and args.mode in ("sketch", "true_topk") and not getattr(args, "replicated_tail", False):
            from .parallel.symm import DistSelect
            self.dsel = DistSelect(peer, min(int(args.k), fm.fp.numel))
            if args.mode == "sketch":
                lo_c, hi_c = DistSelect.coord_range(fm.fp.numel, peer.rank, peer.world)
                self._est_slice = torch.empty(max(hi_c - lo_c, 1), dtype=torch.float32, device=device)

    # -- learning rate -----------------------------------------------------------------
    def get_lr(self):
        """Scalar when there is one parameter group, else a length-d device vector laid
        out in flat-buffer (``model.parameters()``) order -- the reference concatenates per
        *group*, which misaligns it with the gradient for Fixup's three groups (Q9)."""
        groups = self.param_groups
        if len(groups) == 1:
            return groups[0]["lr"]
        key = tuple(g["lr"] for g in groups)
        if key != self._lr_vec_key:
            fp = self.fed_model.fp
            lr_of = {}
            for g in groups:
                for p in g["params"]:
                    lr_of[id(p)] = g["lr"]
            self._lr_vec = fp.segment_vector([lr_of.get(id(p), 0.0) for p in fp.params])
            self._lr_vec_key = key
        return self._lr_vec

    def zero_grad(self, set_to_none=False):
        raise NotImplementedError("call zero_grad() on the model instead")

    # -- the server update ---------------------------------------------------------------
    @torch.no_grad()
    def step(self, closure=None):
        tr = self.fed_model.tracer
        had_round = self.fed_model.pending_samples != 0
        with tr.phase("server"):
            self._step_impl()
        if had_round:
            tr.end_round()

    def _step_impl(self):
        fm, args = self.fed_model, self.args
        lr = self.get_lr()
        if (not torch.is_tensor(lr)) and lr == 0:
            print("WARNING: LR is 0")
        if args.mode == "fedavg":
            if torch.is_tensor(lr):
                raise ValueError("fedavg supports a single learning rate only")
            fm.fedavg_lr = float(lr)
        if fm.pending_samples == 0:
            # nothing new to apply (e.g. the driver's extra step that only publishes the LR
            # for fedavg, cv_train.py:200-203)
            return
        inv_n = 1.0 / fm.pending_samples
        fm.pending_samples = 0
        w = fm.fp.flat
        g = fm.worker.transmit
        rho = args.virtual_momentum
        if fm.reduce_pending:
            fm.reduce_pending = False
            if self._step_fused(inv_n, rho, lr):
                return
        if args.mode == "sketch":
            plan = fm.worker.plan
            if self.verror_buf is None:
                sk.server_update(self.Vvelocity, self.Verror, g, inv_n, rho, args.error_type)
            src = self.Verror if args.error_type != "none" else self.Vvelocity
            if (w.is_cuda and not torch.is_tensor(lr) and not (getattr(args, "p2", 0) and args.p2 > 0)
                    and not getattr(args, "no_cuda_graphs", False)):
                self._sketch_tail_graphed(plan, src, float(lr))
                return
            if getattr(args, "p2", 0) and args.p2 > 0:
                idx, val = self._second_round(plan, src, g, inv_n)
            else:
                idx, val = sk.unsketch_topk(plan, src, args.k)
            sk.apply_zero(plan, idx, val, self.Vvelocity,
                          self.Verror if args.error_type == "virtual" else None, w, lr)
            if args.error_type == "local":
                self.Verror.copy_(self.Vvelocity)
            fm.tracker.record_sparse(idx)
            self.last_update = (idx, val)
        elif args.mode == "true_topk":
            idx, val = dn.true_topk_step(self.Vvelocity, self.Verror, g, inv_n, rho, args.k, w, lr)
            if args.local_momentum > 0:
                fm.worker.mask_velocities(idx, val)
            fm.tracker.record_sparse(idx)
            self.last_update = (idx, val)
        else:
            noise = 0.0
            if args.mode == "uncompressed" and args.do_dp and args.dp_mode == "server":
                noise = args.noise_multiplier
                self._noise_calls += 1
            dn.dense_server_update(self.Vvelocity, g, inv_n, rho, w, 1.0 if args.mode == "fedavg" else lr,
                                   noise_std=noise, seed=args.seed + 7, noise_offset=self._noise_calls << 20)
            fm.tracker.record_dense()

    def _sketch_tail(self, plan, src, lr, record):
        """Heavy-hitter recovery, bucket zeroing + weight update, byte accounting (sketch mode)."""
        fm, args = self.fed_model, self.args
        if self.dsel is not None:
            idx, val = self._unsketch_topk_sharded(plan, src)
        else:
            idx, val = sk.unsketch_topk(plan, src, args.k)
        sk.apply_zero(plan, idx, val, self.Vvelocity,
                      self.Verror if args.error_type == "virtual" else None, fm.fp.flat, lr)
        if args.error_type == "local":
            self.Verror.copy_(self.Vvelocity)
        record(idx)
        return idx, val

    def _unsketch_topk_sharded(self, plan, table):
        """Heavy-hitter recovery with the work divided over the ranks: this rank computes the median estimates of
        its d/W coordinates only (``ceb_sketch_unsketch_est_range``), the exact global top-k comes out of the
        distributed select, and every rank ends up with the same ``(idx, val)`` list."""
        from .ops import native
        from .parallel.symm import DistSelect
        fm = self.fed_model
        peer = fm.rt.peer
        d = fm.fp.numel
        lo, hi = DistSelect.coord_range(d, peer.rank, peer.world)
        scratch = native.select_scratch(fm.device)
        native.call("ceb_sketch_unsketch_est_range", native.ptr(table), native.ptr(plan.packed), plan.n_tokens,
                    native.ptr(plan.blk), plan.num_blocks, plan.r, plan.c, plan.d, lo, hi, native.ptr(self._est_slice),
                    native.ptr(scratch), scratch.numel(), native.stream_ptr(fm.device))
        return self.dsel.run(self._est_slice[:hi - lo], lo, scratch, pass1_done=True)

    def _sketch_tail_graphed(self, plan, src, lr: float):
        """The tail of the sketch-mode server step as ONE CUDA-graph replay (after two eager rounds):
        the learning rate and the round number are read from device memory by the kernels
        (``ceb_sketch_apply_zero_dev`` / ``ceb_tracker_record_dev``), everything else is static."""
        fm = self.fed_model
        tr = fm.tracker
        st = self._tail
        st["calls"] += 1
        if st.get("eager_only") or (st["graph"] is None and st["calls"] <= 2):
            self.last_update = self._sketch_tail(plan, src, lr, tr.record_sparse)
            return
        if st["lr"] is None:
            st["lr"] = torch.zeros(1, dtype=torch.float32, device=fm.device)
        tr.host_advance()                              # may re-allocate the histogram (generation changes)
        if st["graph"] is None or st["gen"] != tr.generation or st["src"] != src.data_ptr():
            tr.round -= 1
            tr.sync_round_dev()                        # device round = rounds applied so far
            tr.round += 1
            # (no eager warm-up on the capture stream: the tail mutates the server state, and it is
            # native kernels only -- nothing in it creates stream-keyed library handles)
            cap = graphs.try_capture(
                fm.device, lambda: self._sketch_tail(plan, src, st["lr"], tr.record_sparse_graph),
                what="sketch server tail", warmup=False)
            if cap is None:
                st["eager_only"] = True
                st["graph"] = None
                tr.round -= 1                          # record_sparse advances the round itself
                self.last_update = self._sketch_tail(plan, src, lr, tr.record_sparse)
                return
            st.update(graph=cap, gen=tr.generation, src=src.data_ptr())
        st["lr"].fill_(lr)
        self.last_update = st["graph"].replay()

    def _step_fused(self, inv_n, rho, lr) -> bool:
        """Peer-memory path: the cross-rank reduce-scatter, the momentum/error (or momentum +
        apply) update of this rank's shard and the all-gather run as ONE kernel over NVLink
        (``csrc/peer.cu``).  Returns True when the whole server step is done."""
        fm, args = self.fed_model, self.args
        peer, w = fm.rt.peer, fm.worker
        lo = self.shard_lo
        vv_shard = self.Vvelocity.view(-1)[lo:]
        if args.mode == "true_topk":
            if self.dsel is None:
                # replicated tail (--replicated_tail): plain all-reduce, then the single-GPU rule on every rank
                peer.all_reduce_buffer_(w.transmit_buf)
                return False
            from .ops import native
            d = fm.fp.numel
            hi = min(self.shard_hi, d)
            ve_shard = self.Verror.view(-1)[lo:]
            peer.fused_topk_state(w.transmit_buf, vv_shard, ve_shard, inv_n, rho)
            scratch = native.select_scratch(fm.device)
            idx, val = self.dsel.run(ve_shard[:max(hi - lo, 0)], lo, scratch, pass1_done=False)
            lr_vec = lr if torch.is_tensor(lr) else None
            native.call("ceb_dense_apply_zero_shard", native.ptr(idx), native.ptr(val), idx.numel(), native.ptr(vv_shard),
                        native.ptr(ve_shard), lo, hi, native.ptr(fm.fp.flat), 0.0 if lr_vec is not None else float(lr),
                        native.ptr(lr_vec), native.stream_ptr(fm.device))
            if args.local_momentum > 0:
                w.mask_velocities(idx, val)
            fm.tracker.record_sparse(idx)
            self.last_update = (idx, val)
            return True
        if args.mode == "sketch":
            ve_shard = self.Verror.view(-1)[lo:]
            peer.fused_sketch(w.transmit_buf, self.verror_buf, vv_shard, ve_shard, inv_n, rho, args.error_type)
            return False        # unsketch / top-k / zeroing / apply follow, replicated on every rank
        peer.fused_dense(w.transmit_buf, w.weights_buf, vv_shard, inv_n, rho,
                         1.0 if args.mode == "fedavg" else lr)
        fm.tracker.record_dense()
        return True

    def _second_round(self, plan, table, g_table, inv_n):
        """FetchSGD second round (``--p2``): recover ``p2*k`` candidates from the sketch,
        fetch the *exact* current-round aggregate at those coordinates from the ranks'
        dense gradient sums, combine it with the sketch estimate of the history part
        (error/momentum carried from earlier rounds), and keep the k largest."""
        fm, args = self.fed_model, self.args
        n_cand = min(plan.d, args.p2 * args.k)
        cand, est = sk.unsketch_topk(plan, table, n_cand)
        dense = getattr(fm.worker, "dense_round_sum", None)
        if dense is None:
            k = min(args.k, n_cand)
            sel, val = topk_sparse(est.contiguous(), k)
            return cand[sel.long()].contiguous(), val
        exact = dense[cand.long()].clone()
        fm.rt.all_reduce_sum_(exact)
        exact *= inv_n
        hist_tab = table - g_table * inv_n
        vals = sk.query(plan, hist_tab, cand) + exact
        sel, val = topk_sparse(vals.contiguous(), min(args.k, n_cand))
        return cand[sel.long()].contiguous(), val


# ----------------------------------------------------------------------------- functional server rules
def args2sketch(args):
    from .csvec import CSVec
    return CSVec(d=args.grad_size, c=args.num_cols, r=args.num_rows, device=args.device,
                 numBlocks=args.num_blocks)


def get_server_update(gradient, Vvelocity, Verror, args, lr):
    """Pure-function form of the server rules with the reference's signature
    (fed_aggregator.py:469-613): returns ``(weight_update, Vvelocity, Verror)`` with a
    *dense* ``weight_update``; ``Vvelocity`` / ``Verror`` are updated in place."""
    mode, rho = args.mode, args.virtual_momentum
    d = args.grad_size
    if mode == "sketch":
        plan = sk.SketchPlan.get(d, args.num_cols, args.num_rows, args.num_blocks, gradient.device)
        sk.server_update(Vvelocity, Verror, gradient.contiguous(), 1.0, rho, args.error_type)
        src = Verror if args.error_type != "none" else Vvelocity
        idx, val = sk.unsketch_topk(plan, src, args.k)
        sk.apply_zero(plan, idx, val, Vvelocity, Verror if args.error_type == "virtual" else None, None, 0.0)
        if args.error_type == "local":
            Verror.copy_(Vvelocity)
        update = torch.zeros(d, dtype=torch.float32, device=gradient.device)
        update[idx.long()] = val
        return update * lr, Vvelocity, Verror
    if mode == "true_topk":
        scratch_w = torch.zeros(d, dtype=torch.float32, device=gradient.device)
        idx, val = dn.true_topk_step(Vvelocity, Verror, gradient.contiguous(), 1.0, rho, args.k, scratch_w, 1.0)
        update = torch.zeros(d, dtype=torch.float32, device=gradient.device)
        update[idx.long()] = val
        return update * lr, Vvelocity, Verror
    if mode in ("uncompressed", "local_topk", "fedavg"):
        Vvelocity.mul_(rho).add_(gradient)
        upd = Vvelocity
        if mode == "uncompressed" and args.do_dp and args.dp_mode == "server":
            upd = Vvelocity + torch.normal(0.0, args.noise_multiplier, size=Vvelocity.shape).to(Vvelocity.device)
        return (upd.clone() if mode == "fedavg" else upd * lr), Vvelocity, Verror
    raise ValueError("unknown mode %r" % (mode,))


def split_results(results, n_results):
    return [np.array([r[i] for r in results]) for i in range(n_results)]
