"""GPT-2 double-heads: model contract and the PersonaChat driver end to end (tiny config, CPU)."""
import os

import torch

from commefficient_b200 import models
from commefficient_b200.gpt2_train import train


def test_gpt2_param_count_matches_reference():
    m = models.GPT2DoubleHeadsModel()
    m.resize_token_embeddings(50257 + 5)
    d = sum(p.numel() for p in m.parameters() if p.requires_grad)
    assert d == 124_444_417          # SURVEY.md section 0 / BASELINE.md
    keys = m.state_dict().keys()
    for k in ("transformer.wte.weight", "transformer.h.0.attn.c_attn.weight", "transformer.h.11.mlp.c_proj.bias",
              "transformer.ln_f.weight", "lm_head.weight", "multiple_choice_head.summary.weight"):
        assert k in keys
    assert m.state_dict()["transformer.h.0.attn.c_attn.weight"].shape == (768, 2304)   # Conv1D (in, out)
    assert m.lm_head.weight.data_ptr() == m.transformer.wte.weight.data_ptr()


def test_gpt2_save_load_roundtrip(tmp_path):
    cfg = models.GPT2Config(vocab_size=100, n_positions=32, n_ctx=32, n_embd=32, n_layer=2, n_head=2)
    m = models.GPT2DoubleHeadsModel(cfg)
    m.save_pretrained(str(tmp_path))
    m2 = models.GPT2DoubleHeadsModel.from_pretrained(str(tmp_path))
    ids = torch.randint(0, 100, (2, 2, 8))
    m.eval(), m2.eval()
    a = m(ids, mc_token_ids=torch.full((2, 2), 7))
    b = m2(ids, mc_token_ids=torch.full((2, 2), 7))
    torch.testing.assert_close(a[0], b[0])
    torch.testing.assert_close(a[1], b[1])


def test_gpt2_driver_end_to_end(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    out = train(["--dataset_name", "PERSONA", "--synthetic", "--dataset_dir", str(tmp_path / "persona"),
                 "--num_clients", "8", "--model_checkpoint", "gpt2-tiny", "--num_workers", "2",
                 "--local_batch_size", "2", "--mode", "sketch", "--error_type", "virtual", "--local_momentum", "0",
                 "--virtual_momentum", "0.9", "--num_epochs", "1", "--k", "2000", "--num_cols", "20000",
                 "--device", "cpu", "--valid_batch_size", "2", "--lr_scale", "0.1"])
    assert out["val_ppl"] < 512 and 0.0 <= out["val_acc"] <= 1.0
    runs = os.listdir(tmp_path / "runs")
    assert runs, "save_pretrained did not write the run directory"


def test_gpt2_client_batching_equals_per_client_rounds():
    """One batched forward/backward over all clients == the per-client loop (loss values, weights)."""
    import numpy as np
    from commefficient_b200.fed_aggregator import FedModel, FedOptimizer
    from commefficient_b200.gpt2_train import compute_loss_train, compute_loss_val
    from commefficient_b200.utils import make_args

    def run(no_batching):
        torch.manual_seed(0)
        cfg = models.GPT2Config(vocab_size=64, n_positions=32, n_ctx=32, n_embd=32, n_layer=1, n_head=2,
                                resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0, summary_first_dropout=0.0)
        model = models.GPT2DoubleHeadsModel(cfg)
        args = make_args(mode="uncompressed", num_workers=3, num_clients=6, local_batch_size=2, device="cpu",
                         dataset_name="PERSONA", local_momentum=0.0, num_results_train=1, weight_decay=0.0,
                         no_client_batching=no_batching, watchdog_timeout=0.0)
        fed = FedModel(model, compute_loss_train, args, compute_loss_val)
        opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args)
        opt.param_groups[0]["lr"] = 0.1
        g = torch.Generator().manual_seed(3)
        ids = torch.randint(0, 64, (6, 2, 12), generator=g)
        labels = ids.clone()
        labels[:, 0] = -1
        labels[0, 1, :8] = -1                       # clients have different numbers of target tokens
        batch = (torch.tensor([0, 0, 1, 1, 2, 2]), ids, torch.full((6, 2), 11), labels,
                 torch.randint(0, 2, (6,), generator=g), ids)
        out = fed(batch)
        opt.step()
        w = fed.fp.flat.clone()
        fed.rt.shutdown()
        return out[0], w

    l_b, w_b = run(False)
    l_c, w_c = run(True)
    np.testing.assert_allclose(l_b, l_c, rtol=1e-5)
    torch.testing.assert_close(w_b, w_c, rtol=1e-4, atol=1e-6)


def _two_rank_gpt2(rank, world, out_dir, argv):
    from commefficient_b200.parallel import runtime
    runtime.init_runtime(rank=rank, world_size=world, device="cpu", backend="gloo")
    os.chdir(out_dir)
    out = train(argv)
    torch.save(out, os.path.join(out_dir, "r%d.pt" % rank))


def test_gpt2_driver_on_two_gloo_ranks(tmp_path):
    """The PersonaChat driver as a two-rank SPMD job: each rank tokenises only its own clients (rank-sharded
    FedSampler), rounds are not skipped, and the validation numbers are those of the one-rank run (no dropout, no
    data augmentation: the two runs see the same gradients)."""
    from commefficient_b200.parallel import runtime
    argv = ["--dataset_name", "PERSONA", "--synthetic", "--dataset_dir", str(tmp_path / "persona"),
            "--num_clients", "8", "--model_checkpoint", "gpt2-tiny", "--num_workers", "2",
            "--local_batch_size", "2", "--mode", "uncompressed", "--error_type", "none", "--local_momentum", "0",
            "--virtual_momentum", "0.9", "--num_epochs", "1", "--device", "cpu", "--valid_batch_size", "2",
            "--lr_scale", "0.1", "--seed", "7"]
    one_dir = tmp_path / "one"
    one_dir.mkdir()
    cwd = os.getcwd()
    os.chdir(one_dir)
    try:
        one = train(argv)
    finally:
        os.chdir(cwd)
    out_dir = tmp_path / "two"
    out_dir.mkdir()
    runtime.spawn(_two_rank_gpt2, 2, args=(str(out_dir), argv), device="cpu")
    r0 = torch.load(out_dir / "r0.pt", weights_only=False)
    r1 = torch.load(out_dir / "r1.pt", weights_only=False)
    assert abs(r0["val_ppl"] - r1["val_ppl"]) < 1e-6
    assert abs(r0["val_ppl"] - one["val_ppl"]) / one["val_ppl"] < 0.05


import pytest  # noqa: E402

_GPT2_BASE = ["--dataset_name", "PERSONA", "--synthetic", "--num_clients", "8", "--model_checkpoint", "gpt2-tiny",
              "--num_workers", "2", "--num_epochs", "1", "--device", "cpu", "--valid_batch_size", "2", "--lr_scale", "0.1",
              "--k", "2000", "--num_cols", "20000"]


@pytest.mark.parametrize("name,flags", [
    ("local_topk", ["--local_batch_size", "2", "--mode", "local_topk", "--error_type", "local", "--local_momentum", "0.9",
                    "--virtual_momentum", "0"]),
    ("true_topk", ["--local_batch_size", "2", "--mode", "true_topk", "--error_type", "virtual", "--local_momentum", "0",
                   "--virtual_momentum", "0.9"]),
    ("fedavg", ["--local_batch_size", "-1", "--mode", "fedavg", "--error_type", "none", "--local_momentum", "0",
                "--fedavg_batch_size", "2", "--num_fedavg_epochs", "1"]),
    ("microbatched_whole_clients", ["--local_batch_size", "-1", "--microbatch_size", "2", "--mode", "sketch", "--error_type",
                                    "virtual", "--local_momentum", "0", "--virtual_momentum", "0.9"]),
    ("four_candidates_two_permutations", ["--local_batch_size", "2", "--num_candidates", "4", "--personality_permutations",
                                          "2", "--mode", "uncompressed", "--error_type", "none", "--local_momentum", "0"]),
    ("test_mode", ["--local_batch_size", "2", "--test", "--mode", "sketch", "--error_type", "virtual", "--local_momentum",
                   "0", "--virtual_momentum", "0.9"]),
])
def test_gpt2_driver_flag_combinations(tmp_path, monkeypatch, name, flags):
    monkeypatch.chdir(tmp_path)
    out = train(_GPT2_BASE + ["--dataset_dir", str(tmp_path / "persona")] + flags)
    assert out["val_ppl"] == out["val_ppl"] and out["val_ppl"] < 1e4 and 0.0 <= out["val_acc"] <= 1.0, (name, out)


def test_gpt2_driver_from_a_checkpoint_directory_with_bpe_files(tmp_path, monkeypatch):
    """--model_checkpoint pointing at a directory in the published GPT-2 layout (config.json, pytorch_model.bin,
    vocab.json, merges.txt): weights load, the byte-level BPE tokenizer is used, the five dialogue tokens are added
    and the embedding matrix grows with them."""
    import json
    from commefficient_b200 import data_utils
    from tests.test_cpu_units import _write_tiny_bpe
    monkeypatch.chdir(tmp_path)
    ck = tmp_path / "gpt2-local"
    corpus = "i like to remodel homes. i like to go hunting. my favorite season is winter. what do you do for a living? " * 4
    _write_tiny_bpe(str(ck), corpus, n_merges=60)
    vocab = json.load(open(ck / "vocab.json"))
    # (a 60-merge vocabulary tokenises almost character by character: sequences run to several hundred tokens)
    cfg = models.GPT2Config(vocab_size=len(vocab), n_positions=1024, n_ctx=1024, n_embd=32, n_layer=2, n_head=2)
    models.GPT2DoubleHeadsModel(cfg).save_pretrained(str(ck))
    assert isinstance(data_utils.load_tokenizer(str(ck)), data_utils.BPETokenizer)
    out = train(["--dataset_name", "PERSONA", "--synthetic", "--dataset_dir", str(tmp_path / "persona"),
                 "--num_clients", "8", "--model_checkpoint", str(ck), "--num_workers", "2", "--local_batch_size", "2",
                 "--mode", "uncompressed", "--error_type", "none", "--local_momentum", "0", "--num_epochs", "1",
                 "--device", "cpu", "--valid_batch_size", "2", "--lr_scale", "0.1"])
    assert out["val_ppl"] == out["val_ppl"] and 0.0 <= out["val_acc"] <= 1.0
    run = sorted((tmp_path / "runs").iterdir())[-1]
    saved = [p.name for p in run.rglob("*") if p.is_file()]
    assert "vocab.json" in saved and "merges.txt" in saved and "added_tokens.json" in saved


def test_gpt2_batched_round_in_client_aligned_chunks(monkeypatch):
    """A batched round larger than MAX_BATCHED_SAMPLES is cut at client boundaries: the per-client token-normalised
    loss (gpt2_train.py:78-97 in the reference) and the weights equal the one-chunk round."""
    import numpy as np
    from commefficient_b200 import fed_worker
    from commefficient_b200.fed_aggregator import FedModel, FedOptimizer
    from commefficient_b200.gpt2_train import compute_loss_train, compute_loss_val
    from commefficient_b200.utils import make_args

    def run(limit):
        monkeypatch.setattr(fed_worker.ClientWorker, "MAX_BATCHED_SAMPLES", limit)
        torch.manual_seed(0)
        cfg = models.GPT2Config(vocab_size=64, n_positions=32, n_ctx=32, n_embd=32, n_layer=1, n_head=2,
                                resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0, summary_first_dropout=0.0)
        model = models.GPT2DoubleHeadsModel(cfg)
        args = make_args(mode="uncompressed", num_workers=3, num_clients=6, local_batch_size=-1, device="cpu",
                         dataset_name="PERSONA", local_momentum=0.0, num_results_train=1, weight_decay=0.0,
                         watchdog_timeout=0.0)
        fed = FedModel(model, compute_loss_train, args, compute_loss_val)
        opt = FedOptimizer(torch.optim.SGD(model.parameters(), lr=1), args)
        opt.param_groups[0]["lr"] = 0.1
        g = torch.Generator().manual_seed(3)
        n = 3 + 2 + 3                                      # three clients with 3, 2 and 3 dialogs
        ids = torch.randint(0, 64, (n, 2, 12), generator=g)
        labels = ids.clone()
        labels[:, 0] = -1
        labels[1, 1, :9] = -1                              # uneven numbers of target tokens inside a client
        labels[6, 1, :4] = -1
        batch = (torch.tensor([0, 0, 0, 1, 1, 2, 2, 2]), ids, torch.full((n, 2), 11), labels,
                 torch.randint(0, 2, (n,), generator=g), ids)
        assert fed.worker.can_batch_clients([3, 2, 3])
        out = fed(batch)
        opt.step()
        w = fed.fp.flat.clone()
        fed.rt.shutdown()
        return np.asarray(out[0]), w

    l_one, w_one = run(4096)
    l_cut, w_cut = run(4)                                  # chunks: clients {0}, {1}, {2} (3 + 2 > 4, 2 + 3 > 4)
    assert fed_worker._client_aligned_chunks((3, 2, 3), 4) == [(0, 3), (3, 5), (5, 8)]
    np.testing.assert_allclose(l_cut, l_one, rtol=1e-5)
    torch.testing.assert_close(w_cut, w_one, rtol=1e-4, atol=1e-6)
