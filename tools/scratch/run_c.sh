#!/bin/bash
export PYTHONPATH=.
timeout 900 python -m pytest tests/test_gpu_e2e.py -x -q 2>&1 | tail -8
for lanes in 1 4 8; do
  echo "lanes=$lanes"
  CEB_CLIENT_LANES=$lanes timeout 300 python bench.py --config resnet18_local_topk --steps 8 --warmup 4 --no_extra 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'])"
  CEB_CLIENT_LANES=$lanes timeout 300 python bench.py --config femnist32_fedavg --steps 6 --warmup 3 --no_extra 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'])"
done
