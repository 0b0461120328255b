#!/usr/bin/env python
"""Entry point with the reference's name: ``python cv_train.py --dataset_name CIFAR10 ...``
(one process = one rank; start N ranks with torchrun for N GPUs)."""
from commefficient_b200.cv_train import main

if __name__ == "__main__":
    main()
