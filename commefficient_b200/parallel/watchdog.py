"""Hang detection and deterministic fault injection.

The reference's only failure handling is queue timeouts: a worker that sees no batch for 120 s
exits silently and the parameter server then blocks in the NCCL reduce or dies with
``queue.Empty`` after 600 s (/root/reference/CommEfficient/fed_worker.py:32-35,
fed_aggregator.py:313-316,360).  Here every rank runs a watchdog
thread: the training loop ``beat()``s at each phase of a round; if no beat arrives within
``--watchdog_timeout`` seconds the rank prints where it was stuck (round, phase, time) and
terminates the whole process with a distinctive exit code, which makes torchrun tear down the
other ranks instead of leaving them spinning on a peer flag or inside a collective.

Fault injection (tests): environment variable ``CEB_FAULT`` =
``hang:<rank>:<round>`` (that rank stops beating and sleeps), ``nan:<rank>:<round>`` (that
rank's transmit buffer is poisoned with NaN), ``kill:<rank>:<round>`` (hard exit).
"""
from __future__ import annotations

import os
import sys
import threading
import time
from typing import Callable, Optional

EXIT_HUNG = 86
EXIT_KILLED = 87


class RoundWatchdog:
    def __init__(self, timeout_s: float, rank: int = 0, on_timeout: Optional[Callable[[str], None]] = None,
                 poll_s: float = 0.25):
        self.timeout_s = float(timeout_s)
        self.rank = rank
        self.on_timeout = on_timeout
        self.poll_s = poll_s
        self._last = time.monotonic()
        self._phase = "init"
        self._round = 0
        self._stop = threading.Event()
        self._armed = False
        self._thread: Optional[threading.Thread] = None

    def start(self):
        if self.timeout_s <= 0 or self._thread is not None:
            return self
        self._thread = threading.Thread(target=self._run, name="ceb-watchdog", daemon=True)
        self._thread.start()
        return self

    def beat(self, phase: str, round_idx: Optional[int] = None):
        self._last = time.monotonic()
        self._phase = phase
        if round_idx is not None:
            self._round = round_idx
        self._armed = True

    def idle(self):
        """Between rounds (data loading, validation by the caller): stop the clock."""
        self._armed = False

    def stop(self):
        self._stop.set()

    def _run(self):
        while not self._stop.wait(self.poll_s):
            if not self._armed:
                continue
            stuck = time.monotonic() - self._last
            if stuck > self.timeout_s:
                msg = ("[commefficient_b200 watchdog] rank %d made no progress for %.1f s "
                       "(round %d, phase %r); aborting all ranks" % (self.rank, stuck, self._round, self._phase))
                print(msg, file=sys.stderr, flush=True)
                if self.on_timeout is not None:
                    self.on_timeout(msg)
                    self._armed = False
                    continue
                os._exit(EXIT_HUNG)


def parse_fault(spec: Optional[str] = None):
    """``(kind, rank, round)`` from ``CEB_FAULT`` or None."""
    spec = spec if spec is not None else os.environ.get("CEB_FAULT", "")
    if not spec:
        return None
    kind, rank, rnd = spec.split(":")
    if kind not in ("hang", "nan", "kill"):
        raise ValueError("CEB_FAULT kind must be hang, nan or kill")
    return kind, int(rank), int(rnd)


def maybe_inject(fault, rank: int, round_idx: int, transmit=None):
    """Apply the configured fault if it targets this rank and round."""
    if fault is None or fault[1] != rank or fault[2] != round_idx:
        return
    kind = fault[0]
    if kind == "kill":
        os._exit(EXIT_KILLED)
    if kind == "nan" and transmit is not None:
        transmit.view(-1)[:1] = float("nan")
    if kind == "hang":
        time.sleep(10 ** 6)
