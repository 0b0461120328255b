#!/usr/bin/env python
"""Entry point with the reference's name: ``python cv_train.py --dataset_name CIFAR10 ...``.  One process = one
rank: start N ranks with torchrun, or pass ``--num_devices N`` and the script starts them itself."""
from commefficient_b200.parallel.runtime import run_driver

if __name__ == "__main__":
    run_driver("commefficient_b200.cv_train", "main")
