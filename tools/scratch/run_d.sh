#!/bin/bash
# launch list of the per-client path (local_topk, ResNet-18): where does a client's 2 ms go?
export PYTHONPATH=.
for lanes in 16; do
  echo "lanes=$lanes"
  timeout 300 python bench.py --config resnet18_local_topk --client_lanes $lanes --steps 8 --warmup 4 --no_extra 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'])"
  timeout 300 python bench.py --config femnist32_fedavg --client_lanes $lanes --steps 6 --warmup 3 --no_extra 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'])"
done
timeout 300 python bench.py --config resnet18_local_topk --client_lanes 1 --no_cuda_graphs --steps 2 --warmup 3 --no_extra --profile > gpurun_out/d_plain.log 2>&1 || { tail -5 gpurun_out/d_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_local_topk.csv python bench.py --config resnet18_local_topk --client_lanes 1 --no_cuda_graphs --steps 2 --warmup 3 --no_extra --profile > gpurun_out/d_ncu.log 2>&1
python tools/summarize_launches.py gpurun_out/launches_local_topk.csv 2 40
