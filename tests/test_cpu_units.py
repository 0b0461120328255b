"""CPU unit tests: CLI, schedules, flat buffers, oracles, sampler, datasets, CSVec API."""
import numpy as np
import pytest
import torch

from commefficient_b200 import data_utils as du
from commefficient_b200 import models
from commefficient_b200.csvec import CSVec
from commefficient_b200.ops import ref
from commefficient_b200.ops.hashing import LARGEPRIME, make_hashes
from commefficient_b200.utils import (FlatParams, PiecewiseLinear, Exp, TableLogger, get_grad, get_param_vec,
                                      make_args, parse_args, set_param_vec, _topk, clip_grad)


def test_cli_defaults_match_reference():
    a = parse_args(argv=["--mode", "uncompressed", "--dataset_name", "CIFAR10"])
    assert (a.k, a.num_cols, a.num_rows, a.num_blocks) == (50000, 500000, 5, 20)
    assert (a.local_momentum, a.virtual_momentum, a.weight_decay, a.num_epochs) == (0.9, 0.0, 5e-4, 24)
    assert (a.port, a.num_workers, a.local_batch_size, a.valid_batch_size, a.seed) == (5315, 1, 8, 8, 21)
    assert a.model == "ResNet9" and a.error_type == "none" and a.max_grad_norm is None
    assert a.do_test is False and a.do_topk_down is False and a.do_dp is False


@pytest.mark.parametrize("bad", [
    dict(mode="fedavg", local_batch_size=8, local_momentum=0.0),
    dict(mode="fedavg", local_batch_size=-1, local_momentum=0.9),
    dict(mode="sketch", local_momentum=0.9, error_type="virtual"),
    dict(mode="sketch", local_momentum=0.0, error_type="none"),
    dict(mode="true_topk", error_type="none"),
    dict(mode="local_topk", error_type="virtual"),
])
def test_cli_rejects_bad_combinations(bad):
    with pytest.raises(ValueError):
        make_args(**bad)


def test_schedules():
    tri = PiecewiseLinear([0, 5, 24], [0, 0.4, 0])
    assert tri(0) == 0 and tri(5) == pytest.approx(0.4) and tri(24) == 0 and tri(2.5) == pytest.approx(0.2)
    assert tri(100) == 0
    e = Exp(2, 1.0, 10)
    assert e(1) == pytest.approx(0.5) and e(12) == pytest.approx(0.1)


def test_flat_params_are_views():
    m = models.ResNet9(num_classes=10, channels={"prep": 4, "layer1": 4, "layer2": 4, "layer3": 4})
    ref_vec = torch.cat([p.data.reshape(-1) for p in m.parameters()])
    fp = FlatParams(m, "cpu")
    m._flat_params = fp
    assert torch.equal(fp.flat, ref_vec)
    fp.flat.mul_(2)
    assert torch.equal(get_param_vec(m), 2 * ref_vec)
    out = m(torch.randn(2, 3, 32, 32)).sum()
    out.backward()
    assert fp.grad.abs().sum() > 0
    args = make_args(mode="uncompressed", weight_decay=0.1, num_workers=4, device="cpu")
    g = get_grad(m, args)
    torch.testing.assert_close(g, fp.grad + 0.025 * fp.flat)
    set_param_vec(m, ref_vec)
    assert torch.equal(next(m.parameters()).data.reshape(-1), ref_vec[: next(m.parameters()).numel()])
    sd = m.state_dict()
    assert "n.prep.conv.weight" in sd and "n.linear.weight" in sd


def test_hashes_follow_csvec_recipe():
    h = make_hashes(1003, 97, 3, 20)
    assert h.n_tokens == 51 and h.buckets.shape == (3, 51) and h.packed.dtype == torch.int32
    assert int(h.buckets.max()) < 97 and int(h.buckets.min()) >= 0
    assert set(h.signs.unique().tolist()) <= {-1.0, 1.0}
    # recompute row 0 by hand with python ints wrapped to int64
    st = torch.random.get_rng_state()
    torch.random.manual_seed(42)
    hh = torch.randint(0, LARGEPRIME, (3, 6), dtype=torch.int64)
    torch.random.set_rng_state(st)

    def wrap(x):
        x &= (1 << 64) - 1
        return x - (1 << 64) if x >= (1 << 63) else x

    for t in (0, 1, 17, 50):
        a = [int(v) for v in hh[0]]
        b = wrap(wrap(a[0] * t) + a[1]) % LARGEPRIME % 97
        assert b == int(h.buckets[0, t])
        s = wrap(wrap(wrap(wrap(wrap(wrap(a[2] * t) + a[3]) * t) + a[4]) * t) + a[5]) % LARGEPRIME % 2 * 2 - 1
        assert s == int(h.signs[0, t])
        packed = int(h.packed[0, t])
        assert (packed & 0x7FFFFFFF) == b and ((packed < 0) == (s < 0))


def test_csvec_api_roundtrip():
    d, c, r = 5000, 2000, 5
    v = torch.zeros(d)
    hot = torch.randperm(d)[:20]
    v[hot] = torch.randn(20) * 10 + 20 * torch.randn(20).sign()
    s = CSVec(d, c, r, device="cpu", numBlocks=4)
    s.accumulateVec(v)
    rec = s.unSketch(k=20)
    assert set(rec.nonzero().view(-1).tolist()) == set(hot.tolist())
    torch.testing.assert_close(rec[hot], v[hot])
    assert s.l2estimate() == pytest.approx(v.norm().item(), rel=0.2)
    s2 = CSVec(d, c, r, device="cpu", numBlocks=4)
    s2.accumulateTable(s.table)
    s2 += s
    torch.testing.assert_close(s2.table, 2 * s.table)
    out = clip_grad(1.0, s2)
    assert out is s2 and s2.l2estimate() == pytest.approx(1.0, rel=1e-4)
    s2.zero()
    assert float(s2.table.abs().sum()) == 0


def test_topk_oracle():
    v = torch.tensor([1.0, -5.0, 3.0, float("nan"), 0.5])
    out = _topk(v, 2)
    assert out.tolist() == [0.0, -5.0, 3.0, 0.0, 0.0]
    m = _topk(torch.tensor([[1.0, 2.0, -3.0], [4.0, -1.0, 0.0]]), 1)
    assert m.tolist() == [[0.0, 0.0, -3.0], [4.0, 0.0, 0.0]]


def test_server_rules_oracle_sketch_zeroes_buckets():
    d, c, r = 2000, 500, 5
    h = make_hashes(d, c, r, 1)
    g = torch.zeros(d)
    g[7] = 10.0
    table = torch.zeros(r, c)
    ref.sketch_accum_ref(table, g, h)
    Vv, Ve = torch.zeros(r, c), torch.zeros(r, c)
    _, (idx, val) = ref.server_update_ref("sketch", table, Vv, Ve, rho=0.9, lr=0.1, k=1, error_type="virtual", hashes=h)
    assert idx.tolist() == [7] and val.item() == pytest.approx(1.0)
    assert float(Ve.abs().sum()) == 0 and float(Vv.abs().sum()) == 0


def test_sampler_and_datasets(tmp_path):
    du.write_synthetic_cifar(str(tmp_path / "c"), images_per_class=20, num_val=30)
    np.random.seed(0)
    ds = du.FedCIFAR10(str(tmp_path / "c"), "CIFAR10", None, False, 40, train=True)
    assert len(ds) == 200 and ds.num_clients == 40 and ds.data_per_client.tolist() == [5] * 40
    cid, img, y = ds[0]
    assert cid == 0 and y == 0
    cid, img, y = ds[199]
    assert cid == 39 and y == 9
    sampler = du.FedSampler(ds, 8, -1, seed=3)
    rounds = list(iter(sampler))
    assert len(rounds) == 5 and all(len(r) == 40 for r in rounds)
    assert sorted(np.concatenate(rounds).tolist()) == list(range(200))
    again = list(iter(du.FedSampler(ds, 8, -1, seed=3)))
    assert all((a == b).all() for a, b in zip(rounds, again))       # every rank derives the same schedule
    iid = du.FedCIFAR10(str(tmp_path / "c"), "CIFAR10", None, True, 7, train=True)
    assert iid.data_per_client.sum() == 200 and iid.num_clients == 7
    val = du.FedCIFAR10(str(tmp_path / "c"), "CIFAR10", None, train=False)
    assert len(val) == 30 and val[3][0] == -1


def test_table_logger_prints(capsys):
    t = TableLogger()
    t.append({"epoch": 1, "lr": 0.1, "acc": np.float32(0.5)})
    t.append({"epoch": 2, "lr": 0.2, "acc": np.float32(0.75)})
    out = capsys.readouterr().out.splitlines()
    assert len(out) == 3 and "epoch" in out[0] and "0.7500" in out[2]


def test_profile_helper_and_shm_helpers(tmp_path):
    import pstats
    from commefficient_b200.utils import profile_helper, shms, sm2np
    out = profile_helper(lambda a, b: a + b, 2, 3, name="add", rank=1, directory=str(tmp_path))
    assert out == 5
    stats = pstats.Stats(str(tmp_path / "add.1.prof"))
    assert stats.total_calls >= 1
    t = torch.arange(6.0).reshape(2, 3)
    v = sm2np(t)
    v[0, 0] = 9.0
    assert float(t[0, 0]) == 9.0 and isinstance(shms(), list)


def test_accumulate_range_cpu_oracle():
    from commefficient_b200.ops import sketch as sk
    plan = sk.SketchPlan.get(1000, 97, 3, 4, "cpu")
    g = torch.randn(1000)
    full = plan.new_table()
    sk.accumulate(plan, full, g)
    parts = plan.new_table()
    for lo, hi in ((0, 10), (10, 600), (600, 1000)):
        sk.accumulate_range(plan, parts, g, lo, hi)
    torch.testing.assert_close(parts, full)


def test_gpt2_regrouped_forward_equals_blockwise_forward():
    """GPT2Model.forward fuses every residual add with the following LayerNorm; it must equal the textbook
    ``for block: x = block(x)`` + ln_f formulation."""
    from commefficient_b200 import models
    import torch.nn.functional as F
    torch.manual_seed(0)
    cfg = models.GPT2Config(vocab_size=97, n_positions=32, n_ctx=32, n_embd=32, n_layer=3, n_head=4,
                            resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0, summary_first_dropout=0.0)
    m = models.GPT2DoubleHeadsModel(cfg).eval()
    tr = m.transformer
    ids = torch.randint(0, 97, (3, 20))
    got = tr(ids, ids)
    x = tr.wte(ids) + tr.wpe(torch.arange(20)) + tr.wte(ids)
    for block in tr.h:
        x = block(x)
    want = F.layer_norm(x, (32,), tr.ln_f.weight, tr.ln_f.bias, tr.ln_f.eps)
    torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-5)


def test_gpt2_label_sparse_loss_equals_full_logits_loss():
    """need_logits=False scores only through lm_rows; on any device it must equal the loss computed
    from the full logits (and the MC loss is untouched)."""
    from commefficient_b200 import models
    cfg = models.GPT2Config(vocab_size=101, n_positions=32, n_ctx=32, n_embd=32, n_layer=2, n_head=4,
                            resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0, summary_first_dropout=0.0)
    m = models.GPT2DoubleHeadsModel(cfg).eval()
    ids = torch.randint(0, 101, (3, 2, 16))
    labels = ids.clone()
    labels[:, 0] = -1
    labels[:, 1, :9] = -1
    mc_tok = torch.full((3, 2), 15)
    mc_lab = torch.tensor([1, 0, 1])
    full = m(ids, token_type_ids=ids, mc_token_ids=mc_tok, lm_labels=labels, mc_labels=mc_lab)
    sparse = m(ids, token_type_ids=ids, mc_token_ids=mc_tok, lm_labels=labels, mc_labels=mc_lab, need_logits=False,
               return_hidden=True)
    assert sparse[2] is None and sparse[-1].shape == (6, 16, 32)
    torch.testing.assert_close(sparse[0], full[0], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(sparse[1], full[1])
    rows, mask = m.lm_rows(sparse[-1], labels)
    assert int(mask.sum()) == 3 * 7 and bool((rows[~mask] == 0).all())      # labels 9..15 -> 7 scored positions


def test_batched_chunks_never_split_a_client():
    """Chunking of a large batched round (ClientWorker.MAX_BATCHED_SAMPLES) keeps every client whole, so losses
    normalised per client (the reference's per-client token mean, gpt2_train.py:78-97) are unaffected."""
    from commefficient_b200.fed_worker import _client_aligned_chunks as chunks
    assert chunks((3, 3, 3, 3), 6) == [(0, 6), (6, 12)]
    assert chunks((10, 2, 2), 4) == [(0, 10), (10, 14)]          # an oversized client stays whole
    assert chunks((5,), 100) == [(0, 5)] and chunks((), 4) == []
    sizes = (7, 1, 4, 4, 9, 2, 2, 2)
    got = chunks(sizes, 8)
    bounds = {0}
    for s in sizes:
        bounds.add(max(bounds) + s)
    assert got[0][0] == 0 and got[-1][1] == sum(sizes)
    assert all(a[1] == b[0] for a, b in zip(got, got[1:]))
    assert all(lo in bounds and hi in bounds for lo, hi in got)


def test_public_header_declares_every_exported_entry_point():
    """include/commefficient_b200.h is the C ABI's contract: every `ceb_*` symbol the library exports is declared
    there (and nothing is declared that the library does not export)."""
    import os
    import re
    import shutil
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib = os.path.join(root, "commefficient_b200", "lib", "libceb200.so")
    if not os.path.exists(lib) or shutil.which("nm") is None:
        pytest.skip("native library or nm not available")
    out = subprocess.run(["nm", "-D", "--defined-only", lib], stdout=subprocess.PIPE, text=True, check=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if ln.split()[-1].startswith("ceb_")}
    header = open(os.path.join(root, "include", "commefficient_b200.h")).read()
    declared = set(re.findall(r"\b(ceb_[a-z0-9_]+)\s*\(", header))
    assert exported - declared == set(), "exported but not declared: %s" % sorted(exported - declared)
    assert declared - exported == set(), "declared but not exported: %s" % sorted(declared - exported)


def test_fed_imagenet_on_a_synthetic_folder_tree(tmp_path):
    """FedImageNet (one client per wnid) over an ImageNet-shaped tree written by write_synthetic_imagenet: natural and
    iid partitions, the ImageNet train / val transforms, and a round through FedSampler + DataLoader."""
    from torch.utils.data import DataLoader
    from commefficient_b200 import data_utils
    root = str(tmp_path)
    wnids = data_utils.write_synthetic_imagenet(root, num_classes=4, images_per_class=3, val_per_class=2)
    assert len(wnids) == 4
    train = data_utils.FedImageNet(root, "ImageNet", data_utils.imagenet_train_transforms, False, 4, train=True, download=False)
    assert len(train) == 12 and train.num_clients == 4 and list(train.data_per_client) == [3, 3, 3, 3]
    cid, img, target = train[7]
    assert int(cid) == 2 and int(target) == 2 and tuple(img.shape) == (3, 224, 224)
    val = data_utils.FedImageNet(root, "ImageNet", data_utils.imagenet_val_transforms, train=False, download=False)
    assert len(val) == 8 and tuple(val[0][1].shape) == (3, 224, 224) and int(val[0][0]) == -1
    iid = data_utils.FedImageNet(root, "ImageNet", data_utils.imagenet_val_transforms, True, 3, train=True, download=False)
    assert iid.num_clients == 3 and sorted(iid.data_per_client) == [4, 4, 4]
    sampler = data_utils.FedSampler(train, num_workers=2, local_batch_size=2, seed=0)
    ids, x, y = next(iter(DataLoader(train, batch_sampler=sampler)))
    assert x.shape == (4, 3, 224, 224) and len(set(ids.tolist())) == 2 and torch.equal(ids, y)


def _write_tiny_bpe(directory, corpus, n_merges=120):
    """vocab.json + merges.txt in GPT-2's format, learned on ``corpus`` by plain pair-frequency BPE."""
    import collections
    import json
    import os
    import regex
    from commefficient_b200.data_utils.fed_persona import _bytes_to_unicode
    table = _bytes_to_unicode()
    pat = regex.compile(r"'s|'t|'re|'ve|'m|'ll|'d| ?\p{L}+| ?\p{N}+| ?[^\s\p{L}\p{N}]+|\s+(?!\S)|\s+")
    words = collections.Counter("".join(table[b] for b in w.encode("utf-8")) for w in pat.findall(corpus))
    splits = {w: list(w) for w in words}
    merges = []
    for _ in range(n_merges):
        pairs = collections.Counter()
        for w, parts in splits.items():
            for a, b in zip(parts, parts[1:]):
                pairs[(a, b)] += words[w]
        if not pairs:
            break
        best = max(sorted(pairs), key=lambda p: pairs[p])
        merges.append(best)
        for w, parts in splits.items():
            out, i = [], 0
            while i < len(parts):
                if i + 1 < len(parts) and (parts[i], parts[i + 1]) == best:
                    out.append(parts[i] + parts[i + 1])
                    i += 2
                else:
                    out.append(parts[i])
                    i += 1
            splits[w] = out
    vocab = {ch: i for i, ch in enumerate(table.values())}
    for a, b in merges:
        vocab.setdefault(a + b, len(vocab))
    vocab["<|endoftext|>"] = len(vocab)
    os.makedirs(directory, exist_ok=True)
    with open(os.path.join(directory, "vocab.json"), "w", encoding="utf-8") as f:
        json.dump(vocab, f, ensure_ascii=False)
    with open(os.path.join(directory, "merges.txt"), "w", encoding="utf-8") as f:
        f.write("#version: 0.2\n" + "\n".join("%s %s" % m for m in merges) + "\n")


def test_bpe_tokenizer_matches_transformers_on_generated_files(tmp_path):
    """BPETokenizer (GPT-2 byte-level BPE over vocab.json + merges.txt) against transformers' GPT2Tokenizer on the
    same files: tokens and ids, including contractions, digits, punctuation runs, repeated spaces and non-ASCII."""
    transformers = pytest.importorskip("transformers")
    from commefficient_b200.data_utils import ATTR_TO_SPECIAL_TOKEN, BPETokenizer, load_tokenizer
    corpus = ("i like to remodel homes. i like to go hunting. i'm a vegan and i don't eat meat! my favorite season is "
              "winter, and my 2 dogs love the snow... what do you do for a living? i work at a café in 1999 - 2003. ") * 3
    d = str(tmp_path / "tok")
    _write_tiny_bpe(d, corpus)
    ours = load_tokenizer(d)
    assert isinstance(ours, BPETokenizer)
    try:
        try:
            hf = transformers.GPT2Tokenizer(vocab=d + "/vocab.json", merges=d + "/merges.txt")      # transformers >= 5
        except TypeError:
            hf = transformers.GPT2Tokenizer(vocab_file=d + "/vocab.json", merges_file=d + "/merges.txt")
    except Exception as exc:          # noqa: BLE001
        pytest.skip("transformers cannot build GPT2Tokenizer from these files here: %s" % exc)
    if not hf.tokenize("i like"):
        pytest.skip("this transformers build ignores the vocabulary files passed to GPT2Tokenizer")
    samples = ["i like to go hunting.", "I'm a vegan, and I don't eat meat!", "my  2 dogs   love the snow...",
               "what do you do for a living?  ", "café naïve — résumé \U0001F600 ok", "1999-2003: we've said 'hello'"]
    for text in samples:
        mine = ours.tokenize(text)
        assert mine == hf.tokenize(text), text
        assert ours.convert_tokens_to_ids(mine) == hf.convert_tokens_to_ids(mine)
    base = len(ours)
    assert ours.add_special_tokens(ATTR_TO_SPECIAL_TOKEN) == 5 and len(ours) == base + 5
    assert ours.add_special_tokens(ATTR_TO_SPECIAL_TOKEN) == 0
    ours.save_pretrained(str(tmp_path / "saved"))
    again = load_tokenizer(str(tmp_path / "saved"))
    assert len(again) == len(ours) and again.tokenize(samples[1]) == ours.tokenize(samples[1])
    assert again.convert_tokens_to_ids("<bos>") == ours.convert_tokens_to_ids("<bos>") >= base


def test_reference_package_name_is_importable():
    """Code written against the reference's package layout imports unchanged."""
    from CommEfficient.fed_aggregator import FedModel, FedOptimizer, get_server_update      # noqa: F401
    from CommEfficient.fed_worker import forward_grad, local_step, process_batch           # noqa: F401
    from CommEfficient.utils import PiecewiseLinear, TableLogger, Timer, parse_args, steps_per_epoch  # noqa: F401
    from CommEfficient.data_utils import FedCIFAR10, FedEMNIST, FedImageNet, FedPERSONA, FedSampler    # noqa: F401
    from CommEfficient.data_utils.fed_sampler import FedSampler as S2
    from CommEfficient.models import FixupResNet50, ResNet9, ResNet18                      # noqa: F401
    import CommEfficient.cv_train as drv
    import commefficient_b200.fed_aggregator as ours
    assert FedModel is ours.FedModel and S2 is FedSampler and hasattr(drv, "main")
