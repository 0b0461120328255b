#!/usr/bin/env python
"""Peer-memory fused collectives vs NCCL (device-timed, max over ranks).

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/bench_collectives.py

Sizes: 5x500 000 sketch table (10 MB), ResNet-9 dense (26.3 MB), GPT-2 dense (497.8 MB).
A two-shot all-reduce brings 2*(N-1)/N * bytes INTO every GPU (its own pulls of the other
ranks' shard copies plus the other ranks' pushes of their reduced shards) and sends the same
amount out; the roofline uses the measured 770 GB/s per direction (B200_PROFILING.md):
ideal = 2*(N-1)/N * bytes / 770e9."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from commefficient_b200.parallel.runtime import init_runtime  # noqa: E402

LINK_GBS = 770.0


def timed(fn, rt, iters=20, warmup=5):
    for _ in range(warmup):
        fn()
    rt.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return rt.max_over_ranks(a.elapsed_time(b)) / iters * 1e3   # us


def main():
    rt = init_runtime(backend="b200", device="cuda")
    peer, dev, W = rt.peer, rt.device, rt.world_size
    rows = []
    for name, n in (("sketch 5x500k", 2_500_000), ("resnet9 dense", 6_568_640), ("gpt2 dense", 124_444_420)):
        buf = peer.symm_empty(peer.padded(n))
        gathered = peer.symm_empty(peer.padded(n))
        buf.tensor.normal_()
        x = torch.randn(n, device=dev)
        lo, hi = peer.shard_range(peer.padded(n))
        Vv = torch.zeros(peer.padded(n) + 4, device=dev)[: peer.padded(n)]
        t_nccl = timed(lambda: dist.all_reduce(x), rt)
        t_peer = timed(lambda: peer.all_reduce_buffer_(buf), rt)
        t_fused = timed(lambda: peer.fused_sketch(buf, gathered, Vv[lo:], gathered.tensor[lo:], 0.01, 0.9, "virtual"), rt)
        variants = {}
        if peer.nvls:      # both data paths on the same buffers: in-switch reduction + multicast vs peer loads / stores
            default = peer.use_nvls
            for name_v, flag in (("nvls", True), ("p2p", False)):
                peer.use_nvls = flag
                variants[name_v] = {
                    "allreduce_us": timed(lambda: peer.all_reduce_buffer_(buf), rt),
                    "fused_us": timed(lambda: peer.fused_sketch(buf, gathered, Vv[lo:], gathered.tensor[lo:], 0.01, 0.9,
                                                                "virtual"), rt)}
            peer.use_nvls = default
        sweep = {}
        if W == 8 and not peer.use_nvls:        # float4s in flight per thread and peer (CEB_PEER_U, read by the library at every launch)
            for u in (1, 2, 4):
                os.environ["CEB_PEER_U"] = str(u)
                sweep["U%d" % u] = timed(lambda: peer.fused_sketch(buf, gathered, Vv[lo:], gathered.tensor[lo:], 0.01, 0.9,
                                                                    "virtual"), rt)
            os.environ.pop("CEB_PEER_U", None)
        # NCCL + separate elementwise server update (what the fused kernel replaces)
        Vv2, Ve2 = torch.zeros(n, device=dev), torch.zeros(n, device=dev)

        def unfused():
            dist.all_reduce(x)
            Vv2.mul_(0.9).add_(x, alpha=0.01)
            Ve2.add_(Vv2)
        t_unfused = timed(unfused, rt)
        ideal = 2 * (W - 1) / W * n * 4 / (LINK_GBS * 1e9) * 1e6
        rows.append({"buffer": name, "bytes": n * 4, "nccl_allreduce_us": t_nccl, "peer_allreduce_us": t_peer,
                     "peer_fused_rs_update_ag_us": t_fused, "nvls_default": bool(peer.use_nvls), "variants": variants,
                     "nccl_plus_update_us": t_unfused,
                     "ideal_us": ideal, "fused_fraction_of_roofline": ideal / t_fused, "fused_us_by_U": sweep})
        for b in (buf, gathered):
            b.close()
            peer.buffers.remove(b)
    if rt.rank == 0:
        print(json.dumps({"n_gpus": W, "link_gbs_measured_ref": LINK_GBS, "rows": rows}, indent=1))
    rt.shutdown()


if __name__ == "__main__":
    main()
