#!/bin/bash
# every BASELINE configuration, ours then the reference, N GPUs (default 1); JSON lines under gpurun_out/
N=${1:-1}
OUT=gpurun_out/bench_all_${N}gpu.jsonl
: > $OUT
run() { if [ "$N" = 1 ]; then python bench.py "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py "$@"; fi; }
for c in resnet9_sketch resnet18_true_topk resnet18_local_topk gpt2_sketch femnist_sketch femnist_uncompressed femnist_fedavg; do
  prec=tf32; [ $c = gpt2_sketch ] && prec=bf16
  run --config $c --gpus $N --steps 8 --warmup 4 --precision $prec 2>gpurun_out/err_$c.log | tail -1 >> $OUT
  if [ -z "$SKIP_REF" ]; then
    run --impl reference --config $c --gpus $N --steps 2 --warmup 1 2>>gpurun_out/err_$c.log | tail -1 >> $OUT
  fi
done
cat $OUT | cut -c1-420
