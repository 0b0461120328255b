"""cv_train.main end to end on the CPU (the reference's driver: /root/reference/CommEfficient/cv_train.py:289-421):
synthetic data in the reference's on-disk formats, a few rounds, validation, the summary."""
import math
import os

import pytest
import torch

from commefficient_b200 import cv_train


def _run(tmp_path, extra):
    argv = ["--synthetic", "--dataset_dir", str(tmp_path / "data"), "--device", "cpu", "--num_epochs", "1",
            "--pivot_epoch", "0.5", "--lr_scale", "0.05", "--num_cols", "2000", "--num_rows", "3", "--k", "300",
            "--valid_batch_size", "8", "--train_dataloader_workers", "0", "--val_dataloader_workers", "0"] + extra
    return cv_train.main(argv)


@pytest.mark.parametrize("mode_flags", [
    ["--mode", "sketch", "--error_type", "virtual", "--local_momentum", "0", "--virtual_momentum", "0.9"],
    ["--mode", "local_topk", "--error_type", "local", "--local_momentum", "0.9", "--virtual_momentum", "0"],
    ["--mode", "fedavg", "--error_type", "none", "--local_momentum", "0", "--num_fedavg_epochs", "1",
     "--fedavg_batch_size", "2", "--local_batch_size", "-1"],
])
def test_cifar10_driver_runs_an_epoch(tmp_path, mode_flags):
    summary = _run(tmp_path, ["--dataset_name", "CIFAR10", "--model", "ResNet9", "--num_clients", "10",
                              "--num_workers", "4", "--local_batch_size", "4"] + mode_flags)
    assert summary is not None
    flat = str(summary)
    assert "nan" not in flat.lower()
    assert os.path.exists(tmp_path / "data" / "stats.json")


def test_imagenet_driver_runs_on_a_synthetic_tree(tmp_path):
    """The imagenet.sh configuration in miniature: FixupResNet50, uncompressed, virtual momentum, one client per class."""
    summary = _run(tmp_path, ["--dataset_name", "ImageNet", "--model", "FixupResNet50", "--num_clients", "4",
                              "--num_workers", "2", "--local_batch_size", "2", "--mode", "uncompressed",
                              "--error_type", "none", "--local_momentum", "0", "--virtual_momentum", "0.9",
                              "--weight_decay", "1e-4"])
    assert summary is not None and "nan" not in str(summary).lower()
    assert os.path.exists(tmp_path / "data" / "meta.bin") and os.path.exists(tmp_path / "data" / "stats.json")


def test_checkpoint_resume_and_finetune_through_the_driver(tmp_path):
    """--checkpoint writes the state_dict and the resume sidecar; --resume continues from it; --finetune starts a
    CIFAR-100 head from the CIFAR-10 checkpoint (reference cv_train.py:336-360)."""
    ck = str(tmp_path / "ck")
    sketch = ["--mode", "sketch", "--error_type", "virtual", "--local_momentum", "0", "--virtual_momentum", "0.9"]
    common = ["--dataset_name", "CIFAR10", "--model", "ResNet9", "--num_clients", "10", "--num_workers", "4",
              "--local_batch_size", "4"] + sketch
    _run(tmp_path, common + ["--checkpoint", "--checkpoint_path", ck])
    assert os.path.exists(os.path.join(ck, "ResNet9.pt")) and os.path.exists(os.path.join(ck, "ResNet9.fedstate.pt"))
    state = torch.load(os.path.join(ck, "ResNet9.pt"))
    assert all(torch.isfinite(v).all() for v in state.values() if v.is_floating_point())
    resumed = _run(tmp_path, common + ["--resume", os.path.join(ck, "ResNet9.fedstate.pt")])
    assert resumed is not None and "nan" not in str(resumed).lower()
    fine = cv_train.main(["--synthetic", "--dataset_dir", str(tmp_path / "data100"), "--device", "cpu", "--num_epochs", "1",
                          "--pivot_epoch", "0.5", "--lr_scale", "0.05", "--num_cols", "2000", "--num_rows", "3", "--k", "300",
                          "--valid_batch_size", "8", "--train_dataloader_workers", "0", "--val_dataloader_workers", "0",
                          "--dataset_name", "CIFAR100", "--model", "ResNet9", "--num_clients", "100", "--num_workers", "4",
                          "--local_batch_size", "2", "--finetune", "--finetuned_from", "CIFAR10", "--finetune_path", ck]
                         + sketch)
    assert fine is not None and "nan" not in str(fine).lower()
