#!/bin/bash
export PYTHONPATH=.
for i in 1 2; do
for f in 1 0; do
  echo "CEB_FUSE_MLP=$f"
  CEB_FUSE_MLP=$f timeout 300 python bench.py --config gpt2_sketch --precision bf16 --steps 20 --warmup 5 --no_extra 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['clocks'])"
done
done
