#!/usr/bin/env bash
# GPT-2 (double heads) on PersonaChat with FetchSGD sketching (BASELINE.json config 4; reference:
# /root/reference/CommEfficient/gpt2_train.py).  --synthetic writes a PersonaChat-shaped dataset
# instead of downloading one.  Usage: scripts/persona_gpt2.sh [NUM_GPUS] [extra flags...]
set -euo pipefail
NGPU="${1:-1}"; shift || true
COMMON=(--dataset_name PERSONA --dataset_dir ./data/persona --model_checkpoint gpt2 --mode sketch
        --num_rows 5 --num_cols 1240000 --k 124000 --error_type virtual --local_momentum 0 --virtual_momentum 0.9
        --num_workers $((8 * NGPU)) --local_batch_size 8 --num_epochs 1 --lr_scale 0.16 --num_candidates 2
        --max_history 2 --defer_results)
if [ "$NGPU" -gt 1 ]; then
  exec python -m torch.distributed.run --nnodes=1 --nproc-per-node "$NGPU" --master-addr 127.0.0.1 \
      gpt2_train.py "${COMMON[@]}" "$@"
fi
exec python gpt2_train.py "${COMMON[@]}" "$@"
