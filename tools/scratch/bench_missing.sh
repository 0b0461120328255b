#!/bin/bash
OUT=gpurun_out/bench_missing_1gpu.jsonl
: > $OUT
timeout 500 python bench.py --impl reference --config gpt2_sketch --gpus 1 --steps 2 --warmup 1 --ref_timeout 450 2>gpurun_out/err_gpt2_ref.log | tail -1 >> $OUT
for c in femnist32_sketch femnist32_uncompressed femnist32_fedavg; do
  timeout 200 python bench.py --config $c --gpus 1 --steps 8 --warmup 4 2>gpurun_out/err_$c.log | tail -1 >> $OUT
  timeout 400 python bench.py --impl reference --config $c --gpus 1 --steps 2 --warmup 1 --ref_timeout 350 2>>gpurun_out/err_$c.log | tail -1 >> $OUT
done
timeout 100 python bench.py --impl reference --config femnist_sketch --gpus 1 --steps 2 --warmup 1 --ref_timeout 90 2>/dev/null | tail -1 >> $OUT
cut -c1-330 $OUT
