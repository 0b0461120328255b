"""CUDA-graph capture that cannot take a round down.

The reference's ``forward_grad`` (/root/reference/CommEfficient/fed_worker.py:249-335) always runs
eagerly; here the launch-bound inner loops (client round, per-client backward, sketch server tail) are
captured and replayed.  Capture is an optimisation, never a requirement:

  * the body is run once EAGERLY on the capture stream first, so library handles / workspaces that
    are keyed by stream (cuDNN, cuBLAS) exist before capture begins -- creating them inside a capture
    is what invalidates it;
  * capture uses ``capture_error_mode="thread_local"`` so unrelated threads (data loaders, the
    watchdog) cannot poison it;
  * any exception during capture is swallowed: the caller gets ``None``, marks the shape "eager only"
    and carries on with the eager result;
  * the per-shape cache is a small LRU, so varying client sizes (non-iid EMNIST / PERSONA) cannot accumulate
    hundreds of graphs with their private memory pools (a pool handle shared across graphs is NOT used: PyTorch
    drops a shared pool when its last graph dies, and the next capture into that handle trips an allocator
    assertion);
  * if a capture does fail after random-number kernels were recorded, the device generator is left with a
    non-zero in-graph offset and every later ``torch.randn`` on the device would raise: the handler rebuilds
    the generator state.
"""
from __future__ import annotations

import warnings
from collections import OrderedDict
from typing import Callable, Optional

import torch

_POOLS = {}
_STREAMS = {}
_WARNED = set()

FORCE_FAIL = False          # tests flip this to exercise the fallback


def capture_stream(device) -> "torch.cuda.Stream":
    key = torch.device(device).index or 0
    s = _STREAMS.get(key)
    if s is None:
        s = torch.cuda.Stream(device=device)
        _STREAMS[key] = s
    return s


def _repair_generator(device) -> None:
    """After a failed capture: give the device's default generator a fresh state object (same seed and offset)
    so that its in-graph bookkeeping is back to "not capturing"."""
    try:
        idx = torch.device(device).index or 0
        gen = torch.cuda.default_generators[idx]
        fresh = torch.Generator(device=torch.device("cuda", idx))
        try:
            fresh.set_state(gen.get_state())
        except Exception:             # noqa: BLE001
            fresh.manual_seed(gen.initial_seed())
        gen.graphsafe_set_state(fresh.graphsafe_get_state())
    except Exception:                 # noqa: BLE001
        pass


class Captured:
    __slots__ = ("graph", "out", "native_launches")

    def __init__(self, graph, out, native_launches):
        self.graph, self.out, self.native_launches = graph, out, native_launches

    def replay(self):
        self.graph.replay()
        if self.native_launches:
            from ..ops import native
            native.add_launch_count(self.native_launches)
        return self.out


def try_capture(device, body: Callable[[], object], what: str = "graph", warmup: bool = True) -> Optional[Captured]:
    """Capture ``body()`` into a CUDA graph.  Returns ``None`` (after a one-time warning) when the
    capture fails for any reason; the device and the capture stream are left usable.  ``body`` must be
    idempotent with respect to the state it reads (it runs once eagerly as warm-up unless ``warmup`` is
    False, then once under capture, which executes nothing)."""
    from ..ops import native
    device = torch.device(device)
    stream = capture_stream(device)
    cur = torch.cuda.current_stream(device)
    try:
        if warmup:
            stream.wait_stream(cur)
            with torch.cuda.stream(stream):
                body()
            cur.wait_stream(stream)
        torch.cuda.synchronize(device)
        before = native.launch_count() if native.available() else 0
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream, capture_error_mode="thread_local"):
            if FORCE_FAIL:
                raise RuntimeError("forced capture failure (test hook)")
            out = body()
        launches = (native.launch_count() - before) if native.available() else 0
        return Captured(graph, out, launches)
    except Exception as exc:          # noqa: BLE001 -- capture is optional by design
        try:
            torch.cuda.synchronize(device)
        except Exception:             # noqa: BLE001
            pass
        _repair_generator(device)
        if what not in _WARNED:
            _WARNED.add(what)
            warnings.warn("CUDA-graph capture of the %s failed (%s: %s); running it eagerly from now on"
                          % (what, type(exc).__name__, str(exc).splitlines()[0] if str(exc) else ""))
        return None


class GraphCache:
    """LRU of per-shape graph entries (``dict`` payloads owned by the caller)."""

    def __init__(self, capacity: int = 12):
        self.capacity = capacity
        self.entries: "OrderedDict[object, dict]" = OrderedDict()

    def get(self, key) -> dict:
        e = self.entries.get(key)
        if e is None:
            e = {"calls": 0, "graph": None, "eager_only": False}
            self.entries[key] = e
            while len(self.entries) > self.capacity:
                self.entries.popitem(last=False)
        else:
            self.entries.move_to_end(key)
        return e

    def __len__(self):
        return len(self.entries)
