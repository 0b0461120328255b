#!/usr/bin/env bash
# ImageNet with FixupResNet50, uncompressed with virtual momentum -- the configuration of the
# reference's imagenet.sh (/root/reference/imagenet.sh:1-21) with its stale flags (--mixup,
# --mixup_alpha, --supervised: not defined by the parser) removed.  One rank per GPU; the reference
# used 7 workers + 1 parameter server on 8 GPUs, here all 8 GPUs compute.
# Usage: scripts/imagenet.sh /path/to/imagenet [extra flags...]
set -euo pipefail
DATA="${1:?path to the ImageNet root}"; shift || true
exec python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 cv_train.py \
    --dataset_name ImageNet --dataset_dir "$DATA" --model FixupResNet50 --mode uncompressed \
    --error_type none --local_momentum 0 --virtual_momentum 0.9 --num_workers 8 --local_batch_size 64 \
    --valid_batch_size 64 --num_epochs 90 --lr_scale 0.1 --weight_decay 1e-4 --k 1000000 --num_rows 1 \
    --num_cols 10000000 --checkpoint "$@"
