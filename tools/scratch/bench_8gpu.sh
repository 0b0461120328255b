#!/bin/bash
N=${1:-8}
OUT=gpurun_out/bench_${N}gpu.jsonl
: > $OUT
tr() { timeout ${T:-240} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@"; }
tr bench.py --gpus $N --steps 20 --warmup 5 2>gpurun_out/err8_a.log | tail -1 >> $OUT
tr bench.py --gpus $N --steps 20 --warmup 5 --no_extra --scaling strong 2>gpurun_out/err8_b.log | tail -1 >> $OUT
tr bench.py --gpus $N --steps 10 --warmup 5 --no_extra --config resnet18_true_topk 2>gpurun_out/err8_c.log | tail -1 >> $OUT
tr bench.py --gpus $N --steps 10 --warmup 5 --no_extra --config gpt2_sketch --precision bf16 2>gpurun_out/err8_d.log | tail -1 >> $OUT
tr bench.py --gpus $N --steps 10 --warmup 5 --no_extra --config femnist_sketch 2>gpurun_out/err8_e.log | tail -1 >> $OUT
tr bench/bench_collectives.py > gpurun_out/collectives_${N}gpu.json 2>gpurun_out/err8_f.log
T=400 tr bench.py --impl reference --gpus $N --steps 2 --warmup 1 --ref_timeout 350 2>gpurun_out/err8_g.log | tail -1 >> $OUT
cut -c1-300 $OUT; tail -c 1500 gpurun_out/collectives_${N}gpu.json
