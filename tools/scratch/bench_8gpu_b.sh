#!/bin/bash
# 8 GPUs, ours only (the reference arm at 8 GPUs is unchanged: profiles/bench_8gpu_ref.json)
N=${1:-8}
export PYTHONPATH=.
OUT=gpurun_out/bench_${N}gpu_r2b.jsonl
: > $OUT
PORT=29700
tr() { PORT=$((PORT + 7)); timeout ${T:-200} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT "$@"; }
tr bench/bench_collectives.py > gpurun_out/collectives_${N}gpu_nvls.json 2>gpurun_out/err8_f.log
tr bench.py --gpus $N --steps 20 --warmup 5 2>gpurun_out/err8_a.log | tail -1 >> $OUT
tr bench.py --gpus $N --steps 10 --warmup 5 --no_extra --config gpt2_sketch --precision bf16 2>gpurun_out/err8_d.log | tail -1 >> $OUT
tr bench.py --gpus $N --steps 8 --warmup 4 --no_extra --config resnet18_local_topk 2>gpurun_out/err8_c.log | tail -1 >> $OUT
tr bench.py --gpus $N --steps 10 --warmup 5 --no_extra --config resnet18_true_topk 2>gpurun_out/err8_e.log | tail -1 >> $OUT
cut -c1-260 $OUT; grep -v "^NCCL" gpurun_out/collectives_${N}gpu_nvls.json | python -c "
import sys, json
d = json.loads(sys.stdin.read())
for r in d['rows']:
    print(r['buffer'], 'nccl', round(r['nccl_allreduce_us'],1), 'fused(default)', round(r['peer_fused_rs_update_ag_us'],1), 'variants', {k: {a: round(b,1) for a,b in v.items()} for k,v in r['variants'].items()}, 'ideal', round(r['ideal_us'],1))
"
