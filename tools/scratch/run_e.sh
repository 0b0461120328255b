#!/bin/bash
export PYTHONPATH=.
timeout 1200 python -m pytest tests/test_gpu_tc.py tests/test_gpu_conv_general.py tests/test_gpu_tf32.py tests/test_gpu_e2e.py -x -q 2>&1 | tail -4
for lanes in 1 4 8; do
  echo "lanes=$lanes"
  for c in resnet18_local_topk femnist32_fedavg femnist_fedavg; do
  timeout 300 python bench.py --config $c --client_lanes $lanes --steps 8 --warmup 4 --no_extra 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['config']['name'], d['value'], d['ms_per_step'], d['e2e']['ms_per_step'])"
  done
done
for c in resnet9_sketch resnet18_true_topk; do
  timeout 300 python bench.py --config $c --steps 20 --warmup 5 --no_extra 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['config']['name'], d['value'], d['ms_per_step'], d['e2e']['ms_per_step'])"
done
