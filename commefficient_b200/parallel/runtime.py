"""SPMD runtime: one rank per GPU (or per CPU process), no parameter-server rank.

The reference is a parameter server + worker processes joined by NCCL ``reduce`` and
host shared memory (/root/reference/CommEfficient/fed_aggregator.py:136-164,301-332,
fed_worker.py:18-35,134-138).  Here every rank is a worker for its share of the
round's clients *and* an aggregator; the only per-round communication is the sum of
the per-rank transmit buffers plus a few hundred bytes of per-client results.

Backends
  ``b200``   peer-memory kernels over NVLink (``parallel.symm`` / ``csrc/peer.cu``)
  ``nccl``   ``torch.distributed`` NCCL -- the baseline the fused path is compared against
  ``gloo``   CPU plumbing / tests (BASELINE.json config 1)
  ``single`` one process, no communication
"""
from __future__ import annotations

import os
from typing import List, Optional

import torch
import torch.distributed as dist

_RUNTIME: Optional["Runtime"] = None


class Runtime:
    def __init__(self, rank: int, world_size: int, device: torch.device, backend: str):
        self.rank = rank
        self.world_size = world_size
        self.device = device
        self.backend = backend
        self.peer = None        # parallel.symm.PeerGroup when backend == "b200"
        self.collective_calls = 0

    @property
    def is_distributed(self) -> bool:
        return self.world_size > 1

    # -- collectives ---------------------------------------------------------------
    def all_reduce_sum_(self, t: torch.Tensor) -> torch.Tensor:
        """In-place sum over ranks; identical result (bitwise) on every rank."""
        if self.world_size == 1:
            return t
        self.collective_calls += 1
        if self.backend == "b200" and self.peer is not None and t.is_cuda:
            return self.peer.all_reduce_sum_(t)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t

    def all_gather_cat(self, t: torch.Tensor, sizes: Optional[List[int]] = None) -> torch.Tensor:
        """Concatenate per-rank 2-D tensors (variable first dimension) on every rank."""
        if self.world_size == 1:
            return t
        self.collective_calls += 1
        if sizes is None:
            n = torch.tensor([t.shape[0]], device=t.device, dtype=torch.int64)
            all_n = [torch.zeros_like(n) for _ in range(self.world_size)]
            dist.all_gather(all_n, n)
            sizes = [int(x.item()) for x in all_n]
        mx = max(sizes) if sizes else 0
        pad = torch.zeros((mx,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad[: t.shape[0]] = t
        outs = [torch.zeros_like(pad) for _ in range(self.world_size)]
        dist.all_gather(outs, pad)
        return torch.cat([o[:s] for o, s in zip(outs, sizes)], dim=0)

    def broadcast_(self, t: torch.Tensor, src: int = 0) -> torch.Tensor:
        if self.world_size > 1:
            dist.broadcast(t, src)
        return t

    def barrier(self):
        if self.world_size > 1:
            if self.device.type == "cuda":
                dist.barrier(device_ids=[self.device.index])
            else:
                dist.barrier()

    def max_over_ranks(self, value: float) -> float:
        if self.world_size == 1:
            return float(value)
        t = torch.tensor([float(value)], dtype=torch.float64,
                         device=self.device if self.device.type == "cuda" else "cpu")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def replicas_agree(self, t: torch.Tensor) -> bool:
        """Whether every rank holds bit-identical ``t`` (a float32 tensor): a 64-bit checksum of the bit
        patterns (sum and position-weighted sum, so permutations show as well) is MIN- and MAX-reduced.
        One pass over ``t``, one tiny collective, one host sync: meant for every N-th round."""
        if self.world_size == 1:
            return True
        flat = t.detach().reshape(-1).view(torch.int32)
        chunk = 1 << 24                                        # bounds the int64 temporaries to 128 MB each
        pos = torch.arange(1, min(chunk, flat.numel()) + 1, device=flat.device, dtype=torch.int64) % 65521
        s0 = torch.zeros((), dtype=torch.int64, device=flat.device)
        s1 = torch.zeros((), dtype=torch.int64, device=flat.device)
        for lo in range(0, flat.numel(), chunk):
            bits = flat[lo:lo + chunk].to(torch.int64)
            s0 += bits.sum()
            s1 += (bits * pos[:bits.numel()]).sum() * (lo // chunk + 1)
        # float64 carries 53 bits exactly: reduce the two 64-bit sums in three 26-bit-wide limbs
        sums = torch.stack([s0, s1])
        limbs = torch.stack([(sums >> s) & 0x3FFFFFF for s in (0, 26, 52)]).reshape(-1).to(torch.float64)
        lo, hi = limbs.clone(), limbs.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        return bool(torch.equal(lo, hi))

    def shutdown(self):
        global _RUNTIME
        if self.peer is not None:
            self.peer.close()
            self.peer = None
        if self.world_size > 1 and dist.is_initialized():
            # leave together: a rank that tears its sockets down while a peer is still finishing its last collective
            # makes that peer's transport throw (gloo: "connection closed by peer") -- seen as a bare non-zero exit
            # (gloo only, and bounded: a peer that already died must not hold the others' exit)
            if self.backend == "gloo":
                try:
                    import datetime
                    dist.monitored_barrier(timeout=datetime.timedelta(seconds=60))
                except Exception:         # noqa: BLE001
                    pass
            try:
                dist.destroy_process_group()
            except Exception:             # noqa: BLE001
                pass
        _RUNTIME = None


def get_runtime() -> Optional[Runtime]:
    return _RUNTIME


def init_runtime(args=None, rank: Optional[int] = None, world_size: Optional[int] = None,
                 backend: Optional[str] = None, device: Optional[str] = None, port: Optional[int] = None,
                 timeout_s: float = 1800.0) -> Runtime:
    """Create (or return) the process-wide runtime.  Rank / world size come from the
    arguments or from the torchrun environment (RANK / WORLD_SIZE / LOCAL_RANK /
    MASTER_ADDR / MASTER_PORT); a plain ``python script.py`` is a 1-rank job."""
    global _RUNTIME
    if _RUNTIME is not None:
        return _RUNTIME
    env = os.environ
    if rank is None:
        rank = int(env.get("RANK", "0"))
    if world_size is None:
        world_size = int(env.get("WORLD_SIZE", "1"))
    dev_kind = device or (getattr(args, "device", None) if args is not None else None)
    if dev_kind is None:
        dev_kind = "cuda" if torch.cuda.is_available() else "cpu"
    if dev_kind.startswith("cuda"):
        local = int(env.get("LOCAL_RANK", str(rank % max(1, torch.cuda.device_count()))))
        torch.cuda.set_device(local)
        dev = torch.device("cuda", local)
    else:
        dev = torch.device("cpu")
    want = backend or (getattr(args, "backend", "auto") if args is not None else "auto")
    if world_size == 1:
        chosen = "single"
    elif want == "auto":
        chosen = "b200" if dev.type == "cuda" else "gloo"
    else:
        chosen = want
    if dev.type == "cpu" and chosen in ("b200", "nccl"):
        chosen = "gloo"
    rt = Runtime(rank, world_size, dev, chosen)
    if world_size > 1 and not dist.is_initialized():
        import datetime
        env.setdefault("MASTER_ADDR", "127.0.0.1")
        if port is None and args is not None and "MASTER_PORT" not in env:
            port = getattr(args, "port", None)
        if port is not None:
            env["MASTER_PORT"] = str(port)
        env.setdefault("MASTER_PORT", "5315")
        dist.init_process_group(
            backend="gloo" if dev.type == "cpu" else "nccl",
            rank=rank, world_size=world_size,
            timeout=datetime.timedelta(seconds=timeout_s),
            **({"device_id": dev} if dev.type == "cuda" else {}))
    if chosen == "b200":
        try:
            from .symm import PeerGroup
            rt.peer = PeerGroup(rt)
        except Exception as exc:  # fall back to NCCL but say so loudly
            # (PeerGroup raises on EVERY rank when any rank fails -- its set-up exchanges a status with
            # each step -- so all ranks take this branch together)
            import warnings
            warnings.warn("peer-memory backend unavailable (%s); using NCCL collectives" % (exc,))
            rt.backend = "nccl"
            rt.peer = None
    _RUNTIME = rt
    return rt


def spawn(fn, nprocs: int, args=(), port: Optional[int] = None, device: str = "cpu"):
    """Run ``fn(rank, world_size, *args)`` in ``nprocs`` processes with the rendezvous
    environment prepared (used by tests and by the drivers for ``--num_devices > 1``
    when not started under torchrun)."""
    import torch.multiprocessing as mp
    from ..utils import free_port
    port = port or free_port(29500 + os.getpid() % 2000)
    ctx = mp.get_context("spawn")
    procs = []
    err_q = ctx.SimpleQueue()
    for rank in range(nprocs):
        p = ctx.Process(target=_spawn_entry, args=(fn, rank, nprocs, port, device, args, err_q))
        p.start()
        procs.append(p)
    for p in procs:
        p.join()
    codes = [p.exitcode for p in procs]
    if any(c != 0 for c in codes):
        msgs = []
        while not err_q.empty():
            msgs.append(err_q.get())
        raise RuntimeError("spawned rank failed (exit codes %s):\n%s" % (codes, "\n".join(msgs)))


def _spawn_entry(fn, rank, world_size, port, device, args, err_q):
    os.environ["RANK"] = str(rank)
    os.environ["WORLD_SIZE"] = str(world_size)
    os.environ["LOCAL_RANK"] = str(rank)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    try:
        fn(rank, world_size, *args)
    except Exception:
        import traceback
        err_q.put("rank %d:\n%s" % (rank, traceback.format_exc()))
        raise


def _driver_rank(rank, world_size, module_name, fn_name, argv):
    import importlib
    getattr(importlib.import_module(module_name), fn_name)(list(argv))


def run_driver(module_name: str, fn_name: str, argv=None):
    """Entry point of the reference-named scripts.  The reference starts its own worker processes from one
    ``python cv_train.py --num_devices N`` invocation (fed_aggregator.py:130-160); here ``--num_devices N`` with N > 1,
    when the process was not started by torchrun (no WORLD_SIZE in the environment), starts N ranks of the same
    program -- one per GPU -- and waits for them.  Under torchrun, or with ``--num_devices 1``, the driver simply runs
    in this process."""
    import importlib
    import sys
    argv = list(sys.argv[1:] if argv is None else argv)
    n = 1
    for i, a in enumerate(argv):
        if a == "--num_devices" and i + 1 < len(argv):
            n = int(argv[i + 1])
        elif a.startswith("--num_devices="):
            n = int(a.split("=", 1)[1])
    if n > 1 and "WORLD_SIZE" not in os.environ:
        device = None
        for i, a in enumerate(argv):
            if a == "--device" and i + 1 < len(argv):
                device = argv[i + 1]
            elif a.startswith("--device="):
                device = a.split("=", 1)[1]
        if device is None:
            device = "cuda" if torch.cuda.is_available() else "cpu"
        if device.startswith("cuda") and torch.cuda.device_count() < n:
            raise SystemExit("--num_devices %d: only %d CUDA devices are visible" % (n, torch.cuda.device_count()))
        spawn(_driver_rank, n, args=(module_name, fn_name, argv), device="cuda" if device.startswith("cuda") else "cpu")
        return None
    return getattr(importlib.import_module(module_name), fn_name)(argv)
