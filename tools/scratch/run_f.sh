#!/bin/bash
# 2 GPUs: NVLS probe + peer kernel tests + collectives bench
export PYTHONPATH=.
python - <<'PY'
import torch
import torch.distributed._symmetric_memory as sm
from torch._C._autograd import DeviceType
print("multicast support:", sm._SymmetricMemory.has_multicast_support(DeviceType.CUDA, 0))
PY
timeout 900 python -m pytest tests/test_distributed.py -x -q -m gpu -s 2>&1 | grep -v "^$" | tail -15
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench/bench_collectives.py 2>&1 | tail -60 > gpurun_out/coll2.json; tail -50 gpurun_out/coll2.json
