"""Parity of the model definition with the upstream modules."""
import pytest
import torch

from conftest import load_reference_module
from fedrec_b200.models import (AdditiveAttention, DistilBertTrunk, MultiHeadAttention, ScaledDotProductAttention,
                                UserEncoder, UserModel)
from fedrec_b200.ops import reference as ref


def test_state_dict_layout_full_size():
    m = UserModel(None, None, "cpu")
    sd = m.state_dict()
    assert len(sd) == 116
    assert sum(v.numel() for v in sd.values()) == 67_527_762
    trainable = [(n, p) for n, p in m.named_parameters() if p.requires_grad]
    assert len(trainable) == 16 and sum(p.numel() for _, p in trainable) == 1_164_882
    expect = {
        "text_encoder.DistillBert.embeddings.word_embeddings.weight": (30522, 768),
        "text_encoder.DistillBert.embeddings.position_embeddings.weight": (512, 768),
        "text_encoder.DistillBert.embeddings.LayerNorm.weight": (768,),
        "text_encoder.DistillBert.transformer.layer.5.attention.out_lin.weight": (768, 768),
        "text_encoder.DistillBert.transformer.layer.0.ffn.lin1.weight": (3072, 768),
        "text_encoder.DistillBert.transformer.layer.3.ffn.lin2.bias": (768,),
        "text_encoder.DistillBert.transformer.layer.2.sa_layer_norm.bias": (768,),
        "text_encoder.DistillBert.transformer.layer.4.output_layer_norm.weight": (768,),
        "text_encoder.additive_attention.att_fc1.weight": (384, 768),
        "text_encoder.additive_attention.att_fc2.weight": (1, 384),
        "text_encoder.fc.weight": (400, 768),
        "user_encoder.multihead_attention.W_Q.weight": (400, 400),
        "user_encoder.multihead_attention.W_V.bias": (400,),
        "user_encoder.additive_attention.att_fc1.weight": (200, 400),
        "user_encoder.additive_attention.att_fc2.bias": (1,),
    }
    for k, shape in expect.items():
        assert tuple(sd[k].shape) == shape, k
    # trainable tensors are views into one flat buffer
    for n, p in trainable:
        assert p.data_ptr() == m.flat.view(n).data_ptr()
        assert p.grad is not None and p.grad.data_ptr() == m.flat.view(n, grad=True).data_ptr()


def test_distilbert_matches_transformers(tiny_bert):
    tr = pytest.importorskip("transformers")
    hf_cfg = tr.DistilBertConfig(vocab_size=tiny_bert.vocab_size, n_layers=tiny_bert.n_layers, n_heads=tiny_bert.n_heads,
                                 dim=tiny_bert.dim, hidden_dim=tiny_bert.hidden_dim)
    hf = tr.DistilBertModel(hf_cfg).eval()
    mine = DistilBertTrunk(tiny_bert).eval()
    assert list(hf.state_dict().keys()) == list(mine.state_dict().keys())
    mine.load_state_dict(hf.state_dict())
    ids = torch.randint(1, 30522, (5, 12))
    mask = torch.ones(5, 12, dtype=torch.long)
    mask[0, 7:] = 0
    mask[3, 3:] = 0
    with torch.no_grad():
        a = hf(ids, attention_mask=mask)[0]
        b = mine(ids, attention_mask=mask)[0]
    torch.testing.assert_close(a, b, atol=2e-5, rtol=1e-4)


def test_attention_modules_match_upstream():
    up = load_reference_module("attention")
    if up is None:
        pytest.skip("upstream sources not available")
    torch.manual_seed(0)
    x = torch.randn(3, 11, 40)
    a_up, a_me = up.AdditiveAttention(40, 16), AdditiveAttention(40, 16)
    a_me.load_state_dict(a_up.state_dict())
    torch.testing.assert_close(a_me(x), a_up(x), atol=1e-6, rtol=1e-5)
    mask = (torch.rand(3, 11) > 0.3).float()
    torch.testing.assert_close(a_me(x, mask), a_up(x, mask), atol=1e-6, rtol=1e-5)
    m_up, m_me = up.MultiHeadAttention(40, 4, 10, 10), MultiHeadAttention(40, 4, 10, 10)
    m_me.load_state_dict(m_up.state_dict())
    torch.testing.assert_close(m_me(x, x, x), m_up(x, x, x), atol=1e-6, rtol=1e-5)
    q = torch.randn(2, 4, 7, 10)
    c1, w1 = ScaledDotProductAttention(10)(q, q, q)
    c2, w2 = up.ScaledDotProductAttention(10)(q, q, q)
    torch.testing.assert_close(c1, c2, atol=1e-6, rtol=1e-5)
    torch.testing.assert_close(w1, w2, atol=1e-6, rtol=1e-5)


def test_stable_normalisation_survives_large_logits():
    from fedrec_b200.models.attention import stable_exp_normalise
    s = torch.tensor([[200.0, 199.0, 0.0]])
    out = stable_exp_normalise(s, -1)
    assert torch.isfinite(out).all() and out.sum() == pytest.approx(1.0, abs=1e-6)
    small = torch.tensor([[-30.0, -31.0]])
    torch.testing.assert_close(stable_exp_normalise(small, -1), ref.exp_normalise(small, -1))  # eps term matters here


def test_user_encoder_functional_reference():
    torch.manual_seed(1)
    ue = UserEncoder().eval()
    x = torch.randn(2, 50, 400) * 0.3
    sd = ue.state_dict()
    got = ref.user_encoder(x, sd["multihead_attention.W_Q.weight"], sd["multihead_attention.W_Q.bias"],
                           sd["multihead_attention.W_K.weight"], sd["multihead_attention.W_K.bias"],
                           sd["multihead_attention.W_V.weight"], sd["multihead_attention.W_V.bias"],
                           sd["additive_attention.att_fc1.weight"], sd["additive_attention.att_fc1.bias"],
                           sd["additive_attention.att_fc2.weight"], sd["additive_attention.att_fc2.bias"], 20)
    torch.testing.assert_close(got, ue(x), atol=1e-5, rtol=1e-4)


def test_pretrained_directory_is_loaded_through_bert_path(tmp_path, tiny_bert):
    """A HuggingFace-format directory (what upstream downloads, encoder.py:19) read from disk: config.json +
    model.safetensors (own reader) or pytorch_model.bin; outputs equal the transformers model's."""
    tr = pytest.importorskip("transformers")
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.models import UserModel
    from fedrec_b200.models.distilbert import read_safetensors
    hf_cfg = tr.DistilBertConfig(vocab_size=500, n_layers=2, n_heads=4, dim=64, hidden_dim=128, max_position_embeddings=40)
    torch.manual_seed(3)
    hf = tr.DistilBertModel(hf_cfg).eval()
    d1 = tmp_path / "st"
    hf.save_pretrained(str(d1))                                   # config.json + model.safetensors
    assert (d1 / "model.safetensors").is_file()
    sd = read_safetensors(str(d1 / "model.safetensors"))
    for k, v in hf.state_dict().items():
        assert torch.equal(sd[k], v), k
    d2 = tmp_path / "bin"
    d2.mkdir()
    (d2 / "config.json").write_text((d1 / "config.json").read_text())
    torch.save({"distilbert." + k: v for k, v in hf.state_dict().items()}, str(d2 / "pytorch_model.bin"))   # MLM-style prefix
    ids = torch.randint(1, 500, (3, 9))
    mask = torch.ones(3, 9, dtype=torch.long)
    mask[1, 5:] = 0
    with torch.no_grad():
        want = hf(ids, attention_mask=mask)[0]
    news_index = torch.zeros(8, 2, 9, dtype=torch.long).numpy()
    for d in (d1, d2):
        trunk = DistilBertTrunk.from_pretrained(str(d)).eval()
        assert trunk.config.n_layers == 2 and trunk.config.dim == 64 and trunk.config.hidden_size == 64
        with torch.no_grad():
            torch.testing.assert_close(trunk(ids, attention_mask=mask)[0], want, atol=2e-5, rtol=1e-4)
        m = UserModel(None, news_index, "cpu", config=FedRecConfig(bert_path=str(d), title_len=9, device="cpu"))
        for k, v in hf.state_dict().items():
            assert torch.equal(m.state_dict()["text_encoder.DistillBert." + k], v), k
        assert not any(p.requires_grad for p in m.text_encoder.DistillBert.parameters())
    # wrong weights for the config are refused, not half-loaded
    bad = tmp_path / "bad"
    bad.mkdir()
    (bad / "config.json").write_text((d1 / "config.json").read_text().replace('"n_layers": 2', '"n_layers": 3'))
    torch.save(hf.state_dict(), str(bad / "pytorch_model.bin"))
    with pytest.raises(ValueError, match="missing"):
        DistilBertTrunk.from_pretrained(str(bad))
    with pytest.raises(FileNotFoundError):
        UserModel(None, news_index, "cpu", config=FedRecConfig(bert_path=str(tmp_path / "nope"), device="cpu"))
