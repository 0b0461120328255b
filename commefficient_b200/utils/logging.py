"""Loggers, timers and run-directory naming.

Behavioural parity with /root/reference/CommEfficient/utils.py:14-99 (Logger,
make_logdir, TableLogger, TSVLogger, Timer, union) plus a JSON-lines perf log
 that the reference does not have.
"""
from __future__ import annotations

import json
import numbers
import os
import time
from datetime import datetime


class Logger:
    """print() shim with the logging-module method names (utils.py:14-24)."""

    def _emit(self, msg, args=None):
        print(msg.format(args))

    debug = info = warn = warning = error = critical = _emit


def union(*dicts):
    out = {}
    for d in dicts:
        out.update(d)
    return out


def make_logdir(args) -> str:
    """``runs/<time>_<W>/<clients>_<mode...>_<k...>`` (utils.py:51-64; the slash
    in ``W/clients`` intentionally yields a nested directory, as upstream)."""
    if args.mode == "sketch":
        mode_str = "%s: %d x %d" % (args.mode, args.num_rows, args.num_cols)
    else:
        mode_str = str(args.mode)
    k_str = "k: %d" % args.k if args.mode in ("sketch", "true_topk", "local_topk") else ""
    stamp = datetime.now().strftime("%b%d_%H-%M-%S")
    name = "_".join([stamp, "%s/%s" % (args.num_workers, args.num_clients), mode_str, k_str])
    return os.path.join("runs", name)


class TableLogger:
    """Fixed-width table; the first row decides the columns (utils.py:66-74)."""

    def __init__(self, stream=None):
        self.stream = stream
        self.keys = None

    def _print(self, *cells):
        print(*cells, file=self.stream)

    def append(self, output):
        if self.keys is None:
            self.keys = list(output.keys())
            self._print(*("{:>12s}".format(k) for k in self.keys))
        cells = []
        for k in self.keys:
            v = output[k]
            if isinstance(v, numbers.Real) and not isinstance(v, numbers.Integral):
                cells.append("{:12.4f}".format(float(v)))
            else:
                cells.append("{:12}".format(v))
        self._print(*cells)


class TSVLogger:
    """DAWNBench submission format (utils.py:76-85)."""

    def __init__(self):
        self.log = ["epoch,hours,top1Accuracy"]

    def append(self, output):
        self.log.append("{},{:.8f},{:.2f}".format(
            output["epoch"], output["total_time"] / 3600.0, output["test_acc"] * 100.0))

    def __str__(self):
        return "\n".join(self.log)


class JSONLLogger:
    """One JSON object per line: per-round perf records (clients/s, phase times)."""

    def __init__(self, path):
        self.path = path
        d = os.path.dirname(path)
        if d:
            os.makedirs(d, exist_ok=True)
        self._f = open(path, "a")

    def append(self, output):
        clean = {}
        for k, v in output.items():
            if hasattr(v, "item"):
                v = v.item()
            clean[k] = v
        self._f.write(json.dumps(clean) + "\n")
        self._f.flush()

    def close(self):
        self._f.close()


class Timer:
    """Wall-clock lap timer (utils.py:89-99)."""

    def __init__(self):
        self.times = [time.time()]
        self.total_time = 0.0

    def __call__(self, include_in_total=True):
        self.times.append(time.time())
        dt = self.times[-1] - self.times[-2]
        if include_in_total:
            self.total_time += dt
        return dt


class DeviceTimer:
    """CUDA-event phase timer.  ``with t.phase("sketch"):`` records a pair of
    events on the current stream; ``summary()`` synchronises once and returns
    per-phase milliseconds.  On CPU it falls back to ``time.perf_counter``.
    (No reference counterpart; SURVEY.md section 5.1 "Ours".)"""

    def __init__(self, enabled=True, device="cuda"):
        import torch
        self.enabled = enabled
        self.cuda = enabled and device != "cpu" and torch.cuda.is_available()
        self._open = []
        self._records = []

    class _Phase:
        def __init__(self, owner, name):
            self.owner, self.name = owner, name

        def __enter__(self):
            o = self.owner
            if not o.enabled:
                return self
            if o.cuda:
                import torch
                self.start = torch.cuda.Event(enable_timing=True)
                self.stop = torch.cuda.Event(enable_timing=True)
                self.start.record()
            else:
                self.start = time.perf_counter()
            return self

        def __exit__(self, *exc):
            o = self.owner
            if not o.enabled:
                return False
            if o.cuda:
                self.stop.record()
                o._records.append((self.name, self.start, self.stop))
            else:
                o._records.append((self.name, self.start, time.perf_counter()))
            return False

    def phase(self, name):
        return DeviceTimer._Phase(self, name)

    def summary(self, reset=True):
        out = {}
        if self.cuda and self._records:
            import torch
            torch.cuda.synchronize()
        for name, a, b in self._records:
            ms = a.elapsed_time(b) if self.cuda else (b - a) * 1e3
            out[name] = out.get(name, 0.0) + ms
        if reset:
            self._records = []
        return out


class PhaseTracer:
    """Per-round phase tracing (SURVEY.md section 5.1 "Ours"; the reference only has a wall-clock
    ``Timer``).  Every phase is an NVTX range plus a CUDA-event pair on the current stream; every
    ``every`` rounds the accumulated per-phase milliseconds are reduced to the MAX over ranks
    (``reduce_max``: a callable taking/returning a list of floats) and rank 0 appends one JSON line
    ``{"round": n, "rounds": m, "ms_per_round": {phase: ms}}`` to ``path``.  Disabled (zero
    overhead beyond an attribute test) when ``path`` is empty."""

    def __init__(self, path, every=50, rank=0, reduce_max=None, device="cuda"):
        self.enabled = bool(path)
        self.path, self.every, self.rank, self.reduce_max = path, max(1, int(every)), rank, reduce_max
        self.timer = DeviceTimer(self.enabled, device)
        self.rounds_open = 0
        self.round_no = 0
        self._nvtx = None
        if self.enabled and self.timer.cuda:
            import torch
            self._nvtx = torch.cuda.nvtx

    class _Range:
        def __init__(self, owner, name):
            self.owner, self.name = owner, name
            self.inner = owner.timer.phase(name)

        def __enter__(self):
            if self.owner._nvtx is not None:
                self.owner._nvtx.range_push(self.name)
            self.inner.__enter__()
            return self

        def __exit__(self, *exc):
            self.inner.__exit__(*exc)
            if self.owner._nvtx is not None:
                self.owner._nvtx.range_pop()
            return False

    class _Null:
        def __enter__(self):
            return self

        def __exit__(self, *exc):
            return False

    _NULL = _Null()

    def phase(self, name):
        if not self.enabled:
            return PhaseTracer._NULL
        return PhaseTracer._Range(self, name)

    def end_round(self):
        """Call once per federated round (after the server step)."""
        if not self.enabled:
            return None
        self.round_no += 1
        self.rounds_open += 1
        if self.rounds_open < self.every:
            return None
        return self.flush()

    def flush(self):
        if not self.enabled or self.rounds_open == 0:
            return None
        summ = self.timer.summary()
        names = sorted(summ)
        vals = [summ[n] / self.rounds_open for n in names]
        if self.reduce_max is not None:
            vals = self.reduce_max(vals)
        rec = {"round": self.round_no, "rounds": self.rounds_open,
               "ms_per_round": {n: round(float(v), 4) for n, v in zip(names, vals)}}
        self.rounds_open = 0
        if self.rank == 0:
            with open(self.path, "a") as f:
                f.write(json.dumps(rec) + "\n")
        return rec


def profile_helper(fn, *args, name="round", rank=0, directory="profile", **kwargs):
    """Run ``fn(*args, **kwargs)`` under cProfile and dump ``<directory>/<name>.<rank>.prof`` (host-side
    hot spots of the Python orchestration).  Reference: ``fed_aggregator.py:46-52`` defines the same
    hook around a function that no longer exists and never calls it; here it is usable:
    ``profile_helper(model, batch)`` profiles one round.  Device time belongs to ``PhaseTracer`` / ncu."""
    import cProfile
    os.makedirs(directory, exist_ok=True)
    prof = cProfile.Profile()
    try:
        return prof.runcall(fn, *args, **kwargs)
    finally:
        prof.dump_stats(os.path.join(directory, "%s.%d.prof" % (name, rank)))


def sm2np(tensor):
    """numpy view of a CPU (e.g. shared-memory or pinned) tensor without a copy (reference:
    ``utils.py:299-303``, unused there)."""
    return tensor.detach().numpy()


def shms():
    """Names under /dev/shm -- the reference's leak check for its shared-memory client state
    (``fed_aggregator.py:23-25``).  This framework keeps client state in device / pinned memory and
    creates no shm segments of its own, so the list should not grow with ``num_clients``."""
    try:
        return sorted(os.listdir("/dev/shm"))
    except OSError:
        return []
