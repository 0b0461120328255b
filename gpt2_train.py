#!/usr/bin/env python
"""Entry point with the reference's name: ``python gpt2_train.py --dataset_name PERSONA ...``."""
from commefficient_b200.gpt2_train import train

if __name__ == "__main__":
    train()
