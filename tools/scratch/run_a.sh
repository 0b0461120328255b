#!/bin/bash
# LN forward vec + fp32 segment BN + GELU-in-epilogue: tests and the two affected benches
export PYTHONPATH=.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc.py tests/test_gpu_attn.py tests/test_gpu_tf32.py tests/test_gpu_e2e.py -x -q 2>&1 | tail -8
for prec in bf16 tf32; do
  timeout 300 python bench.py --config gpt2_sketch --precision $prec --steps 10 --warmup 4 --no_extra 2>&1 | tail -1
  timeout 300 python bench.py --config resnet18_true_topk --precision $prec --steps 10 --warmup 4 --no_extra 2>&1 | tail -1
done
