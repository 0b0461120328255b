#!/bin/bash
# every BASELINE configuration on one GPU, ours only (reference arms: profiles/bench_configs_1gpu.jsonl)
export PYTHONPATH=.
OUT=gpurun_out/bench_all_1gpu_r2b.jsonl
: > $OUT
for c in resnet9_sketch resnet18_true_topk resnet18_local_topk gpt2_sketch femnist_sketch femnist_uncompressed femnist_fedavg femnist32_sketch femnist32_uncompressed femnist32_fedavg; do
  prec=tf32; [ $c = gpt2_sketch ] && prec=bf16
  extra="--no_extra"; [ $c = resnet9_sketch ] && extra=""
  timeout 200 python bench.py --config $c --gpus 1 --steps 10 --warmup 4 --precision $prec $extra 2>gpurun_out/err_$c.log | tail -1 >> $OUT
done
python - <<'PY'
import json
for l in open("gpurun_out/bench_all_1gpu_r2b.jsonl"):
    d = json.loads(l)
    print(d["config"]["name"], d["dtype"], round(d["value"], 1), round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), d["gpu_launches"], d.get("bf16", {}).get("value"))
PY
