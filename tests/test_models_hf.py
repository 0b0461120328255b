"""The transformer trunks against the installed ``transformers`` implementations (random init, copied weights):
GPT-2 (pre-LN, final LayerNorm) and GPT-1 (post-LN, none) -- the two classes the reference picks between at
/root/reference/CommEfficient/gpt2_train.py:263-268."""
import pytest
import torch

from commefficient_b200.models.gpt2 import GPT2Config, GPT2Model, OpenAIGPTDoubleHeadsModel

transformers = pytest.importorskip("transformers")


@pytest.mark.parametrize("post_ln", [False, True])
def test_trunk_matches_transformers(post_ln):
    torch.manual_seed(0)
    kw = dict(vocab_size=97, n_positions=32, n_embd=64, n_layer=2, n_head=4, resid_pdrop=0.0, embd_pdrop=0.0,
              attn_pdrop=0.0)
    if post_ln:
        hf = transformers.OpenAIGPTModel(transformers.OpenAIGPTConfig(afn="gelu", **kw)).eval()
    else:
        hf = transformers.GPT2Model(transformers.GPT2Config(**kw)).eval()
    ours = GPT2Model(GPT2Config(vocab_size=97, n_positions=32, n_ctx=32, n_embd=64, n_layer=2, n_head=4,
                                post_ln=post_ln)).eval()
    mine = ours.state_dict()
    renamed = {}
    for k, v in hf.state_dict().items():
        k2 = k.replace("tokens_embed", "wte").replace("positions_embed", "wpe")
        if k2 in mine and mine[k2].shape == v.shape:
            renamed[k2] = v
    assert sorted(renamed) == sorted(mine), "every parameter of ours has a counterpart of the same shape"
    ours.load_state_dict(renamed)
    ids = torch.randint(0, 97, (3, 17))
    types = torch.randint(0, 97, (3, 17))
    with torch.no_grad():
        want = hf(input_ids=ids, token_type_ids=types).last_hidden_state
        got = ours(ids, types)
    torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-5)


def test_gpt1_class_is_post_ln_without_final_layernorm():
    cfg = GPT2Config(vocab_size=50, n_positions=16, n_ctx=16, n_embd=32, n_layer=1, n_head=2)
    m = OpenAIGPTDoubleHeadsModel(cfg)
    assert m.config.post_ln and not hasattr(m.transformer, "ln_f")
    ids = torch.randint(0, 50, (2, 2, 9))
    out = m(ids, token_type_ids=ids, mc_token_ids=torch.full((2, 2), 8), lm_labels=ids, mc_labels=torch.zeros(2, dtype=torch.long))
    assert torch.isfinite(out[0]) and torch.isfinite(out[1])


def test_double_heads_model_matches_transformers():
    """GPT2DoubleHeadsModel: LM loss, multiple-choice loss and both logits against transformers' class of the same
    name with copied weights (labels: -1 here = -100 there; the reference calls the pytorch_transformers ancestor of
    that class, /root/reference/CommEfficient/gpt2_train.py:76-99)."""
    from commefficient_b200.models.gpt2 import GPT2DoubleHeadsModel
    torch.manual_seed(1)
    kw = dict(vocab_size=101, n_positions=32, n_embd=64, n_layer=2, n_head=4, resid_pdrop=0.0, embd_pdrop=0.0,
              attn_pdrop=0.0)
    hf = transformers.GPT2DoubleHeadsModel(transformers.GPT2Config(summary_first_dropout=0.0, num_labels=1, **kw)).eval()
    ours = GPT2DoubleHeadsModel(GPT2Config(vocab_size=101, n_positions=32, n_ctx=32, n_embd=64, n_layer=2, n_head=4,
                                           summary_first_dropout=0.0)).eval()
    mine = ours.state_dict()
    src = hf.state_dict()
    loaded = {k: src[k] for k in mine if k in src and src[k].shape == mine[k].shape}
    assert sorted(loaded) == sorted(mine), sorted(set(mine) - set(loaded))
    ours.load_state_dict(loaded)
    B, C, T = 3, 2, 13
    ids = torch.randint(0, 101, (B, C, T))
    types = torch.randint(0, 101, (B, C, T))
    mc_ids = torch.randint(0, T, (B, C))
    lm = torch.randint(0, 101, (B, C, T))
    lm[:, 0] = -1                                   # only the last candidate carries LM labels
    lm[:, :, :4] = -1
    mc = torch.randint(0, C, (B,))
    with torch.no_grad():
        want = hf(input_ids=ids, token_type_ids=types, mc_token_ids=mc_ids, labels=lm.masked_fill(lm == -1, -100),
                  mc_labels=mc)
        lm_loss, mc_loss, lm_logits, mc_logits = ours(ids, token_type_ids=types, mc_token_ids=mc_ids, lm_labels=lm,
                                                      mc_labels=mc)
    torch.testing.assert_close(lm_logits, want.logits, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(mc_logits, want.mc_logits, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(lm_loss, want.loss, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(mc_loss, want.mc_loss, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("hf_cls", ["GPT2DoubleHeadsModel", "GPT2LMHeadModel", "GPT2Model"])
def test_from_pretrained_reads_transformers_checkpoints(tmp_path, hf_cls):
    """A directory written by transformers' ``save_pretrained`` (safetensors; with or without the ``transformer.``
    prefix; with or without the heads) loads, and the trunk computes what the saved model computes."""
    from commefficient_b200.models.gpt2 import GPT2DoubleHeadsModel
    torch.manual_seed(2)
    kw = dict(vocab_size=101, n_positions=32, n_embd=64, n_layer=2, n_head=4, resid_pdrop=0.0, embd_pdrop=0.0,
              attn_pdrop=0.0)
    hf = getattr(transformers, hf_cls)(transformers.GPT2Config(**kw)).eval()
    hf.save_pretrained(str(tmp_path))
    ours = GPT2DoubleHeadsModel.from_pretrained(str(tmp_path)).eval()
    ids = torch.randint(0, 101, (3, 11))
    with torch.no_grad():
        base = hf.transformer if hasattr(hf, "transformer") else hf
        want = base(input_ids=ids).last_hidden_state
        got = ours.transformer(ids)
    torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-5)
    assert ours.lm_head.weight.data_ptr() == ours.transformer.wte.weight.data_ptr()


def test_from_pretrained_refuses_a_checkpoint_without_the_trunk(tmp_path):
    from commefficient_b200.models.gpt2 import GPT2DoubleHeadsModel
    cfg = GPT2Config(vocab_size=50, n_positions=16, n_ctx=16, n_embd=32, n_layer=1, n_head=2)
    cfg.to_json_file(str(tmp_path / "config.json"))
    torch.save({"unrelated.weight": torch.zeros(3)}, str(tmp_path / "pytorch_model.bin"))
    with pytest.raises(KeyError):
        GPT2DoubleHeadsModel.from_pretrained(str(tmp_path))


def test_gpt1_from_pretrained_reads_a_transformers_checkpoint(tmp_path):
    torch.manual_seed(3)
    kw = dict(vocab_size=97, n_positions=32, n_embd=64, n_layer=2, n_head=4, resid_pdrop=0.0, embd_pdrop=0.0,
              attn_pdrop=0.0)
    hf = transformers.OpenAIGPTLMHeadModel(transformers.OpenAIGPTConfig(afn="gelu", **kw)).eval()
    hf.save_pretrained(str(tmp_path))
    ours = OpenAIGPTDoubleHeadsModel.from_pretrained(str(tmp_path)).eval()
    assert ours.config.post_ln and not hasattr(ours.transformer, "ln_f")
    ids = torch.randint(0, 97, (2, 9))
    with torch.no_grad():
        want = hf.transformer(input_ids=ids).last_hidden_state
        got = ours.transformer(ids)
    torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-5)
