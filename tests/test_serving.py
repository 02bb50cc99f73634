"""engine/serving.py: top-k recommendation from a resident news-vector table gives the scores and the order of
UserModel.forward (model.py:118-119) over the whole catalogue."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedrec_b200.config import FedRecConfig  # noqa: E402
from fedrec_b200.engine.serving import Recommender  # noqa: E402
from fedrec_b200.models import UserModel  # noqa: E402


@pytest.fixture()
def served(tiny_bert, tiny_data):
    torch.manual_seed(0)
    cfg = FedRecConfig(title_len=12, device="cpu", max_his_len=50)
    m = UserModel(None, tiny_data.news_index, "cpu", config=cfg, bert_config=tiny_bert).eval()
    return m, Recommender(m).refresh()


def test_scores_are_the_models(served):
    m, rec = served
    N = m.news_tokens.shape[0]
    g = torch.Generator().manual_seed(1)
    his = torch.randint(1, N, (6, 50), generator=g)
    his[0, 20:] = 0
    cand = torch.randint(1, N, (6, 7), generator=g)
    with torch.no_grad():
        want = m(cand, his, compute_loss=False)[0]
    torch.testing.assert_close(rec.score(his, cand), want, atol=1e-5, rtol=1e-4)


def test_top_k_is_the_brute_force_ranking(served):
    m, rec = served
    N = m.news_tokens.shape[0]
    his = rec.pad_histories([[5, 9, 11], list(range(1, 80)), []])
    assert his.shape == (3, 50) and his[0, :3].tolist() == [5, 9, 11] and his[1, 0] == 30 and his[2].sum() == 0
    everything = torch.arange(N).repeat(3, 1)
    full = rec.score(his, everything)
    idx, sc = rec.recommend(his, k=8, exclude_seen=False)
    full[:, 0] = -1.0
    want = torch.topk(full, 8, dim=1)
    assert torch.equal(idx, want.indices)
    torch.testing.assert_close(sc, want.values, atol=1e-6, rtol=1e-5)
    # seen items and the padding row never come back; an allow-list restricts the catalogue
    idx2, _ = rec.recommend(his, k=8, exclude_seen=True)
    for b in range(3):
        assert not (set(idx2[b].tolist()) & (set(his[b].tolist()) | {0}))
    allowed = torch.zeros(N, dtype=torch.bool)
    allowed[100:120] = True
    idx3, _ = rec.recommend(his, k=5, allowed=allowed)
    assert ((idx3 >= 100) & (idx3 < 120)).all()


def test_refresh_follows_the_model_and_checks_sizes(served):
    m, rec = served
    v0, ver = rec.news_vecs.clone(), rec.version
    with torch.no_grad():
        m.text_encoder.fc.weight.mul_(1.5)
    rec.refresh()
    assert rec.version == ver + 1 and not torch.allclose(rec.news_vecs, v0)
    with pytest.raises(ValueError):
        rec.refresh(torch.zeros(3, 400))
    half = Recommender(m, table_dtype=torch.bfloat16).refresh()
    his = rec.pad_histories([[5, 9, 11]])
    a, _ = rec.recommend(his, k=20)
    b, _ = half.recommend(his, k=20)
    assert len(set(a[0].tolist()) & set(b[0].tolist())) >= 15          # a bf16 table keeps the ranking up to near-ties


@pytest.mark.gpu
def test_on_the_gpu_the_table_comes_from_the_sm100_kernels(tiny_data):
    """DistilBERT-base shapes so the fused encoder is the one that runs; vectors within bf16 tolerance of the fp32 torch
    path, top-k consistent with the scores."""
    from fedrec_b200.data import make_synthetic_mind
    from fedrec_b200.ops import dispatch as D
    torch.manual_seed(0)
    dev = torch.device("cuda", 0)
    data = make_synthetic_mind(num_news=700, num_train=8, num_valid=2, title_len=30, seed=1)
    m = UserModel(None, data.news_index, dev, config=FedRecConfig(title_len=30, num_news=700)).eval()
    n0 = D.launch_count()
    rec = Recommender(m, encode_batch=512).refresh()
    assert D.launch_count() > n0, "the news table was not encoded by the native kernels"
    ids = torch.arange(700, device=dev)
    ref = m.encode_unique(ids, keep_cache=False, use_fused=False)
    err = (rec.news_vecs - ref).norm() / ref.norm()
    assert err < 3e-2, float(err)
    his = rec.pad_histories([[5, 9, 11], list(range(1, 80))]).to(dev)
    idx, sc = rec.recommend(his, k=10)
    full = rec.score(his, ids.repeat(2, 1))
    full[:, 0] = -1.0
    full.scatter_(1, his, -1.0)
    want = torch.topk(full, 10, dim=1)
    torch.testing.assert_close(sc, want.values, atol=1e-5, rtol=1e-4)
    assert idx.shape == (2, 10) and not (set(idx[1].tolist()) & set(range(30, 80)))


def test_json_lines_front_end(tmp_path, tiny_data):
    import io
    import json
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts"))
    import recommend
    cfg = FedRecConfig(synthetic=True, num_news=200, title_len=8, device="cpu", batch_size=2, bert_layers=1, bert_dim=64,
                       bert_heads=4, bert_ffn=128, snapshot_path="")
    d = __import__("fedrec_b200.data", fromlist=["make_synthetic_mind"]).make_synthetic_mind(200, 8, 2, 8, 50, seed=cfg.seed)
    nids = [n for n in d.nid2index if n != "<unk>"]
    reqs = [{"user": "a", "history": nids[:3], "k": 4}, {"user": "b", "history": ["nope"], "k": 2}, {"user": "c", "history": nids[5:40]}]
    out = io.StringIO()
    n = recommend.serve([json.dumps(r) for r in reqs] + [""], out, cfg, default_k=3)
    rows = [json.loads(l) for l in out.getvalue().splitlines()]
    assert n == 3 and [r["user"] for r in rows] == ["a", "b", "c"] and [len(r["news"]) for r in rows] == [4, 2, 3]
    assert not set(rows[0]["news"]) & set(nids[:3]) and "<unk>" not in rows[1]["news"]
    assert all(rows[i]["scores"] == sorted(rows[i]["scores"], reverse=True) for i in range(3))


def test_the_table_round_trips_through_a_file(served, tmp_path):
    m, rec = served
    p = str(tmp_path / "news_vecs.pt")
    rec.save_table(p)
    other = Recommender(m).load_table(p)
    assert other.version == rec.version and torch.equal(other.news_vecs, rec.news_vecs)
    his = rec.pad_histories([[3, 4, 5]])
    assert torch.equal(other.recommend(his, k=5)[0], rec.recommend(his, k=5)[0])
