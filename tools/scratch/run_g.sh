#!/bin/bash
# 2 GPUs with the NVLS kernels forced on: the whole multi-GPU test tier, then a lanes bench under the peer backend
export PYTHONPATH=.
CEB_NVLS_MIN_WORLD=2 timeout 900 python -m pytest tests/test_distributed.py -x -q -m gpu 2>&1 | grep -v "^$\|NCCL version" | tail -6
tr() { timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@"; }
CEB_NVLS_MIN_WORLD=2 tr bench.py --gpus 2 --steps 10 --warmup 5 --no_extra 2>gpurun_out/g_a.log | tail -1 | cut -c1-200
tr bench.py --gpus 2 --steps 10 --warmup 5 --no_extra 2>gpurun_out/g_b.log | tail -1 | cut -c1-200
tr bench.py --gpus 2 --steps 8 --warmup 4 --no_extra --config resnet18_local_topk 2>gpurun_out/g_c.log | tail -1 | cut -c1-200
