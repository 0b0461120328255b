#!/usr/bin/env python
"""Entry point with the reference's name: ``python gpt2_train.py --dataset_name PERSONA ...`` (``--num_devices N``
starts N ranks, one per GPU, unless torchrun already did)."""
from commefficient_b200.parallel.runtime import run_driver

if __name__ == "__main__":
    run_driver("commefficient_b200.gpt2_train", "train")
