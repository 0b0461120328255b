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
