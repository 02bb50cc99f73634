"""DistilBERT trunk / text-head kernels vs the fp32 nn.Module code."""
import pytest
import torch
import torch.nn.functional as F

from fedrec_b200 import ops
from fedrec_b200.config import FedRecConfig
from fedrec_b200.data import make_synthetic_mind
from fedrec_b200.models import UserModel
from fedrec_b200.ops import dispatch as D

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def model():
    ops.require_native()
    torch.manual_seed(0)
    data = make_synthetic_mind(num_news=400, num_train=8, num_valid=2, title_len=30, seed=1)
    m = UserModel(None, data.news_index, DEV, config=FedRecConfig(title_len=30))
    # random-init LayerNorm/bias values are trivial (1/0): perturb so the test exercises them
    with torch.no_grad():
        for n, p in m.named_parameters():
            if "LayerNorm" in n or "layer_norm" in n or n.endswith(".bias"):
                p.add_(torch.randn_like(p) * 0.1)
    return m


def test_layernorm(model):
    x = torch.randn(1000, 768, device=DEV).bfloat16()
    g, b = torch.randn(768, device=DEV), torch.randn(768, device=DEV)
    out = D.layernorm_bf16(x, g, b)
    want = F.layer_norm(x.float(), (768,), g, b, 1e-12)
    torch.testing.assert_close(out.float(), want, atol=3e-2, rtol=2e-2)
    cnt = torch.tensor([10], device=DEV, dtype=torch.int32)
    out2 = torch.zeros_like(x)
    D.layernorm_bf16(x, g, b, out=out2, count=cnt, rows_per_item=30)
    assert (out2[300:] == 0).all() and torch.equal(out2[:300], out[:300])


def test_embed_ln(model):
    bert = model.text_encoder.DistillBert.eval()
    from fedrec_b200.engine.fused import PackedTrunk
    tr = PackedTrunk(bert)
    items = torch.tensor([5, 1, 399, 0, 17], device=DEV, dtype=torch.int32)
    out = torch.zeros(5 * 30, 768, device=DEV, dtype=torch.bfloat16)
    D.bert_embed_ln(model.news_tokens, items, None, 5, 30, tr.word, tr.pos, tr.emb_g, tr.emb_b, out)
    with torch.no_grad():
        want = bert.embeddings(model.news_tokens[items.long()].long()).view(-1, 768)
    torch.testing.assert_close(out.float(), want, atol=4e-2, rtol=3e-2)


@pytest.mark.parametrize("impl", ["tc", "simt"])
@pytest.mark.parametrize("n,first", [(9, 1), (301, 0), (4, 0)])
def test_attention(model, impl, n, first):
    torch.manual_seed(3)
    T, H = 30, 12
    qkv = (torch.randn(n * T, 2304, device=DEV) * 0.7).bfloat16()
    items = torch.arange(first, first + n, device=DEV, dtype=torch.int32)      # first=0 includes the all-masked <unk> row
    out = torch.zeros(n * T, 768, device=DEV, dtype=torch.bfloat16)
    D.bert_attention(qkv, model.news_mask, items, None, n, T, H, out, impl=impl)
    q, k, v = [t.float().view(n, T, H, 64).transpose(1, 2) for t in qkv.split(768, dim=1)]
    mask = model.news_mask[items.long()]
    s = (q / 8.0) @ k.transpose(-1, -2)
    s = s.masked_fill((mask == 0)[:, None, None, :], torch.finfo(torch.float32).min)
    want = (torch.softmax(s, -1) @ v).transpose(1, 2).reshape(n * T, 768)
    torch.testing.assert_close(out.float(), want, atol=2e-2, rtol=2e-2)
    if n > 100:                                                                # device-side count: rows past it stay untouched
        cnt = torch.tensor([77], device=DEV, dtype=torch.int32)
        out2 = torch.zeros_like(out)
        D.bert_attention(qkv, model.news_mask, items, cnt, n, T, H, out2, impl=impl)
        assert torch.equal(out2[: 77 * T], out[: 77 * T]) and (out2[77 * T:] == 0).all()


def test_attention_other_title_lengths():
    torch.manual_seed(4)
    for T in (50, 64, 17):
        n, H = 21, 12
        qkv = (torch.randn(n * T, 2304, device=DEV) * 0.7).bfloat16()
        mask = (torch.rand(n, T, device=DEV) > 0.3).int()
        mask[:, 0] = 1
        outs = []
        for impl in ("tc", "simt"):
            out = torch.zeros(n * T, 768, device=DEV, dtype=torch.bfloat16)
            D.bert_attention(qkv, mask, None, None, n, T, H, out, impl=impl)
            outs.append(out.float())
        q, k, v = [t.float().view(n, T, H, 64).transpose(1, 2) for t in qkv.split(768, dim=1)]
        s = ((q / 8.0) @ k.transpose(-1, -2)).masked_fill((mask == 0)[:, None, None, :], torch.finfo(torch.float32).min)
        want = (torch.softmax(s, -1) @ v).transpose(1, 2).reshape(n * T, 768)
        for o in outs:
            torch.testing.assert_close(o, want, atol=2e-2, rtol=2e-2)


def test_trunk_and_head_match_module(model):
    from fedrec_b200.engine.fused import FusedStep
    fs = FusedStep(model, model.cfg)
    uniq = torch.tensor([3, 7, 11, 200, 399, 1, 0, 0], device=DEV, dtype=torch.int32)
    count = torch.tensor([6], device=DEV, dtype=torch.int32)
    b = fs.encode(uniq, count, 8)
    torch.cuda.synchronize()
    te = model.text_encoder.eval()
    with torch.no_grad():
        ids = uniq[:6].long()
        hidden = te.trunk(model.news_tokens[ids].long(), model.news_mask[ids])
        vec = te.head(hidden, None)
    got_h = b["x"][: 6 * 30].float().view(6, 30, 768)
    # bf16 activations through 6 layers: compare with a relative Frobenius criterion and a loose elementwise one
    rel = float((got_h - hidden).norm() / hidden.norm())
    assert rel < 2e-2, rel
    relv = float((b["news_vecs"][:6] - vec).norm() / vec.norm())
    assert relv < 2e-2, relv
    torch.testing.assert_close(b["news_vecs"][:6], vec, atol=3e-2, rtol=5e-2)
