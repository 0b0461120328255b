#!/bin/bash
export PYTHONPATH=.
tr() { timeout 110 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@"; }
echo "== lanes 8"
CEB_HANG_DUMP=70 tr bench.py --gpus 2 --steps 6 --warmup 4 --no_extra --config resnet18_local_topk > gpurun_out/h_a.log 2>&1; tail -1 gpurun_out/h_a.log | cut -c1-250
grep -n "File \"/root/repo\|Thread\|Error\|error" gpurun_out/h_a.log | head -40
echo "== lanes 1"
CEB_HANG_DUMP=70 tr bench.py --gpus 2 --steps 6 --warmup 4 --no_extra --config resnet18_local_topk --client_lanes 1 > gpurun_out/h_b.log 2>&1; tail -1 gpurun_out/h_b.log | cut -c1-250
grep -n "File \"/root/repo\|Thread\|Error\|error" gpurun_out/h_b.log | head -40
