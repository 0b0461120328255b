#!/bin/bash
# last GPU call of the round: smoke(), the default bench, and the clipped / DP variants of the headline mode
export PYTHONPATH=.
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 200 python bench.py 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('default', d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d.get('bf16',{}).get('value'))"
timeout 200 python bench.py --no_extra --steps 6 --warmup 4 --set max_grad_norm=1.0 2>gpurun_out/err_clip.log | tail -1 | tee gpurun_out/bench_clip.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('clipped sketch', d['value'], d['ms_per_step'], d['config'].get('client_batching'))"
timeout 200 python bench.py --no_extra --steps 6 --warmup 4 --set max_grad_norm=1.0 --client_lanes 1 2>>gpurun_out/err_clip.log | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('clipped sketch, 1 lane', d['value'], d['ms_per_step'])"
