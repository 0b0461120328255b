#!/usr/bin/env bash
# FetchSGD on CIFAR-10 (the headline configuration of BASELINE.md): 10 000 clients x 5 images, 100
# clients per round per GPU, 5 x 500 000 sketch, k = 50 000, virtual momentum + virtual error.
# Usage: scripts/cifar10_fetchsgd.sh [NUM_GPUS] [extra flags...]     (--synthetic: no dataset download)
set -euo pipefail
NGPU="${1:-1}"; shift || true
COMMON=(--dataset_name CIFAR10 --dataset_dir ./data/cifar10 --model ResNet9 --mode sketch
        --num_rows 5 --num_cols 500000 --k 50000 --error_type virtual --local_momentum 0 --virtual_momentum 0.9
        --num_clients 10000 --num_workers $((100 * NGPU)) --local_batch_size -1 --num_epochs 24 --lr_scale 0.4
        --pivot_epoch 5 --weight_decay 5e-4 --defer_results)
if [ "$NGPU" -gt 1 ]; then
  exec python -m torch.distributed.run --nnodes=1 --nproc-per-node "$NGPU" --master-addr 127.0.0.1 \
      cv_train.py "${COMMON[@]}" "$@"
fi
exec python cv_train.py "${COMMON[@]}" "$@"
