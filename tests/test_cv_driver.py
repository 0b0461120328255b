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


def _two_rank_driver(rank, world, out_dir, argv):
    from commefficient_b200.parallel import runtime
    runtime.init_runtime(rank=rank, world_size=world, device="cpu", backend="gloo")
    summary = cv_train.main(argv)
    torch.save({"summary": summary}, os.path.join(out_dir, "r%d.pt" % rank))


def test_driver_on_two_gloo_ranks_matches_one_rank(tmp_path):
    """The whole driver as an SPMD job (two gloo ranks, rank-sharded data loading through attach_sampler; the replica
    checksum runs every second round) ends close to where the one-rank run ends: not bit-equal, because the random
    crops / flips of the training transform are drawn per loaded sample, and each rank now loads only its own."""
    from commefficient_b200.parallel import runtime
    argv = ["--synthetic", "--dataset_dir", str(tmp_path / "data"), "--device", "cpu", "--num_epochs", "1",
            "--pivot_epoch", "0.5", "--lr_scale", "0.05", "--num_cols", "2000", "--num_rows", "3", "--k", "300",
            "--valid_batch_size", "8", "--train_dataloader_workers", "0", "--val_dataloader_workers", "0",
            "--dataset_name", "CIFAR10", "--model", "ResNet9", "--num_clients", "10", "--num_workers", "4",
            "--local_batch_size", "4", "--mode", "sketch", "--error_type", "virtual", "--local_momentum", "0",
            "--virtual_momentum", "0.9", "--checkpoint"]
    one = str(tmp_path / "one")
    one_summary = cv_train.main(argv + ["--checkpoint_path", one])
    two = str(tmp_path / "two")
    out_dir = str(tmp_path / "out")
    os.makedirs(out_dir)
    runtime.spawn(_two_rank_driver, 2, args=(out_dir, argv + ["--checkpoint_path", two, "--replica_check_every", "2"]),
                  device="cpu")
    w1 = torch.load(os.path.join(one, "ResNet9.pt"))
    w2 = torch.load(os.path.join(two, "ResNet9.pt"))
    assert os.path.exists(os.path.join(two, "ResNet9.fedstate.pt.rank0")) and os.path.exists(
        os.path.join(two, "ResNet9.fedstate.pt.rank1"))
    num = sum(float((w1[k].float() - w2[k].float()).pow(2).sum()) for k in w1 if w1[k].is_floating_point())
    den = sum(float(w1[k].float().pow(2).sum()) for k in w1 if w1[k].is_floating_point())
    assert math.sqrt(num / den) < 1e-2
    # validation: every rank evaluated only its own shards (ShardedEvalSampler) and all ranks report the whole set
    two_summary = [torch.load(os.path.join(out_dir, "r%d.pt" % r), weights_only=False)["summary"] for r in range(2)]
    assert float(two_summary[0]["test_loss"]) == float(two_summary[1]["test_loss"])
    assert abs(float(two_summary[0]["test_loss"]) - float(one_summary["test_loss"])) < 1e-2
    assert abs(float(two_summary[0]["test_acc"]) - float(one_summary["test_acc"])) < 0.05


_SKETCH = ["--mode", "sketch", "--error_type", "virtual", "--local_momentum", "0", "--virtual_momentum", "0.9"]
_C10 = ["--dataset_name", "CIFAR10", "--num_clients", "10", "--local_batch_size", "4", "--model", "ResNet9", "--num_workers", "4"]


@pytest.mark.parametrize("name,flags", [
    ("whole_client_batches", _C10 + ["--local_batch_size", "-1"] + _SKETCH),
    ("true_topk_local_momentum", _C10 + ["--mode", "true_topk", "--error_type", "virtual", "--local_momentum", "0.9",
                                         "--virtual_momentum", "0"]),
    ("topk_down", _C10 + ["--mode", "true_topk", "--error_type", "virtual", "--local_momentum", "0",
                          "--virtual_momentum", "0.9", "--topk_down"]),
    ("dp_worker", _C10 + ["--mode", "uncompressed", "--error_type", "none", "--local_momentum", "0", "--dp",
                          "--dp_mode", "worker", "--noise_multiplier", "0.01", "--l2_norm_clip", "1.0"]),
    ("dp_server", _C10 + ["--mode", "uncompressed", "--error_type", "none", "--local_momentum", "0", "--dp",
                          "--dp_mode", "server", "--noise_multiplier", "0.01"]),
    ("iid_eval_before_start", _C10 + ["--iid", "--eval_before_start"] + _SKETCH),
    ("emnist_whole_clients", ["--dataset_name", "EMNIST", "--num_clients", "12", "--local_batch_size", "-1", "--model",
                              "ResNet9", "--num_workers", "4"] + _SKETCH),
    ("fixup_resnet9_clipped", _C10 + ["--model", "FixupResNet9", "--max_grad_norm", "1.0"] + _SKETCH),
    ("microbatches", _C10 + ["--microbatch_size", "2", "--mode", "uncompressed", "--error_type", "none",
                             "--local_momentum", "0"]),
    ("batchnorm_resnet9", _C10 + ["--batchnorm"] + _SKETCH),
    ("test_mode", _C10 + ["--test"] + _SKETCH),
])
def test_driver_flag_combinations(tmp_path, name, flags):
    """One epoch through the driver for the flag combinations the reference's README / scripts use: nothing raises,
    the summary row is finite."""
    summary = _run(tmp_path, flags)
    assert summary is not None, name
    for key in ("train_loss", "test_loss", "test_acc"):
        assert math.isfinite(float(summary[key])), (name, key, summary[key])


def test_driver_two_ranks_with_per_client_state(tmp_path):
    """local_topk with local momentum / error on two gloo ranks through the driver: clients are assigned to the rank
    that owns their state rows (id mod 2), the sampler shards by the same rule, whole-client batches; the replicas
    stay identical (checked every round) and the resume sidecars hold each rank's rows."""
    from commefficient_b200.parallel import runtime
    argv = ["--synthetic", "--dataset_dir", str(tmp_path / "data"), "--device", "cpu", "--num_epochs", "1",
            "--pivot_epoch", "0.5", "--lr_scale", "0.05", "--k", "300", "--valid_batch_size", "8",
            "--train_dataloader_workers", "0", "--val_dataloader_workers", "0", "--dataset_name", "CIFAR10", "--model",
            "ResNet9", "--num_clients", "10", "--num_workers", "4", "--local_batch_size", "-1", "--mode", "local_topk",
            "--error_type", "local", "--local_momentum", "0.9", "--virtual_momentum", "0", "--checkpoint",
            "--checkpoint_path", str(tmp_path / "ck"), "--replica_check_every", "1"]
    out_dir = str(tmp_path / "out")
    os.makedirs(out_dir)
    runtime.spawn(_two_rank_driver, 2, args=(out_dir, argv), device="cpu")
    s0 = torch.load(os.path.join(out_dir, "r0.pt"), weights_only=False)["summary"]
    s1 = torch.load(os.path.join(out_dir, "r1.pt"), weights_only=False)["summary"]
    assert s0 is not None and math.isfinite(float(s0["train_loss"]))
    assert float(s0["train_loss"]) == float(s1["train_loss"]) and float(s0["test_acc"]) == float(s1["test_acc"])
    side = [torch.load(os.path.join(str(tmp_path / "ck"), "ResNet9.fedstate.pt.rank%d" % r), weights_only=False)
            for r in range(2)]
    owned = [set(int(c) for c in s["client_velocities"].keys()) for s in side]
    assert owned[0] and owned[1] and all(c % 2 == 0 for c in owned[0]) and all(c % 2 == 1 for c in owned[1])
    assert [set(int(c) for c in s["client_errors"].keys()) for s in side] == owned


def test_num_devices_flag_starts_that_many_ranks(tmp_path):
    """``python cv_train.py --num_devices 2`` (the reference's way of asking for two devices) starts two ranks of the
    driver by itself; rank 0 prints the table once and both ranks write their resume sidecars."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ck = tmp_path / "ck"
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_PORT")}
    res = subprocess.run(
        [sys.executable, os.path.join(root, "cv_train.py"), "--num_devices", "2", "--device", "cpu", "--synthetic",
         "--dataset_dir", str(tmp_path / "data"), "--dataset_name", "CIFAR10", "--model", "ResNet9", "--num_clients", "10",
         "--num_workers", "4", "--local_batch_size", "4", "--num_epochs", "1", "--pivot_epoch", "0.5", "--lr_scale", "0.05",
         "--num_cols", "2000", "--num_rows", "3", "--k", "300", "--valid_batch_size", "8", "--mode", "sketch",
         "--error_type", "virtual", "--local_momentum", "0", "--virtual_momentum", "0.9", "--checkpoint",
         "--checkpoint_path", str(ck)],
        cwd=str(tmp_path), env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    assert res.stdout.count("Total Upload (MiB)") == 1
    assert (ck / "ResNet9.fedstate.pt.rank0").exists() and (ck / "ResNet9.fedstate.pt.rank1").exists()
