import numpy as np
import pytest
import torch

from conftest import load_reference_module
from fedrec_b200.utils import metrics as M


def test_golden_values():
    # values computed from the upstream evaluation_functions.py during the survey
    a = M.compute_amn([1, 0, 0, 1, 0], [.1, .4, .35, .8, .2])
    assert a == pytest.approx((0.5, 0.6, 0.850345, 0.850345), abs=1e-5)
    b = M.compute_amn([1, 0, 0, 0, 0], [.55, .6, .5, .52, .51])
    assert b == pytest.approx((0.75, 0.5, 0.630930, 0.630930), abs=1e-5)


def test_against_upstream_module():
    up = load_reference_module("evaluation_functions")
    if up is None:
        pytest.skip("upstream sources not available")
    rng = np.random.default_rng(0)
    for _ in range(50):
        n = int(rng.integers(3, 40))
        y = (rng.random(n) < 0.3).astype(int)
        if y.sum() in (0, n):
            y[0], y[1] = 1, 0
        s = rng.random(n)
        assert M.compute_amn(y, s) == pytest.approx(up.compute_amn(y, s), abs=1e-9)


def test_auc_ties_match_sklearn():
    from sklearn.metrics import roc_auc_score
    y = [1, 0, 1, 0, 0, 1]
    s = [0.5, 0.5, 0.7, 0.1, 0.7, 0.2]
    assert M.roc_auc_score(y, s) == pytest.approx(roc_auc_score(y, s))


def test_batched_matches_numpy():
    rng = np.random.default_rng(1)
    S, C = 64, 5
    scores = rng.random((S, C)).astype(np.float32)
    labels = np.zeros((S, C), dtype=np.float32)
    labels[:, 0] = 1
    out = M.batched_ranking_metrics(torch.tensor(scores), torch.tensor(labels))
    for i in range(S):
        a = M.compute_amn(labels[i], scores[i])
        got = (out["auc"][i].item(), out["mrr"][i].item(), out["ndcg5"][i].item(), out["ndcg10"][i].item())
        assert got == pytest.approx(a, abs=1e-5)


def test_evaluation_split():
    rng = np.random.default_rng(2)
    nid2index = {f"N{i}": i for i in range(20)}
    news_vecs = rng.standard_normal((20, 8))
    user_vecs = rng.standard_normal((3, 8))
    samples = [(0, ["N1", "N2"], ["N3", "N4", "N5"], [], "u"), (1, ["N6"], ["N7", "N8"], [], "u"),
               (2, ["N9"], ["N10"], [], "u")]
    out = M.evaluation_split(news_vecs, user_vecs, samples, nid2index)
    assert out.shape == (3, 4)
    up = load_reference_module("evaluation_functions")
    if up is not None:
        np.testing.assert_allclose(out, up.evaluation_split(news_vecs, user_vecs, samples, nid2index), atol=1e-9)


def test_tied_scores_rank_like_the_reference():
    """All-equal scores with the positive in column 0: upstream's argsort()[::-1] puts it last."""
    import torch
    from fedrec_b200.utils.metrics import batched_ranking_metrics, compute_amn
    y = [1, 0, 0, 0, 0]
    s = [0.5] * 5
    auc, mrr, n5, n10 = compute_amn(y, s)
    m = batched_ranking_metrics(torch.tensor([s]), torch.tensor([y]))
    assert float(m["mrr"][0]) == pytest.approx(mrr) == pytest.approx(0.2)
    assert float(m["ndcg5"][0]) == pytest.approx(n5) and float(m["ndcg10"][0]) == pytest.approx(n10)
    assert float(m["auc"][0]) == pytest.approx(auc) == pytest.approx(0.5)
    # partial ties
    y2, s2 = [1, 0, 0, 1, 0], [0.7, 0.7, 0.2, 0.2, 0.9]
    ref = compute_amn(y2, s2)
    m2 = batched_ranking_metrics(torch.tensor([s2]), torch.tensor([y2]))
    assert [float(m2[k][0]) for k in ("auc", "mrr", "ndcg5", "ndcg10")] == pytest.approx(list(ref))


def test_ragged_rows_match_the_per_impression_numpy_metrics():
    """``valid`` mask: impressions with different candidate counts padded to one width give, row by row, what
    ``compute_amn`` gives on the unpadded lists (ties and several positives included)."""
    import numpy as np
    import torch
    from fedrec_b200.utils.metrics import batched_ranking_metrics, compute_amn
    rng = np.random.default_rng(0)
    rows = []
    for n in (2, 3, 7, 12, 30, 5):
        y = np.zeros(n, dtype=np.int64)
        y[rng.choice(n, size=rng.integers(1, max(2, n // 3)), replace=False)] = 1
        if y.all():
            y[0] = 0
        s = np.round(rng.random(n), 1)                     # coarse scores: plenty of ties
        rows.append((y, s))
    C = max(len(y) for y, _ in rows)
    scores = torch.full((len(rows), C), 0.77)              # padding carries a real-looking score on purpose
    labels = torch.ones(len(rows), C)                      # ... and a positive label: the mask must neutralise both
    valid = torch.zeros(len(rows), C, dtype=torch.bool)
    for i, (y, s) in enumerate(rows):
        scores[i, : len(y)] = torch.from_numpy(s).float()
        labels[i, : len(y)] = torch.from_numpy(y).float()
        valid[i, : len(y)] = True
    m = batched_ranking_metrics(scores, labels, valid)
    for i, (y, s) in enumerate(rows):
        want = compute_amn(y, s.astype(np.float32))
        got = [float(m[k][i]) for k in ("auc", "mrr", "ndcg5", "ndcg10")]
        assert np.allclose(got, want, atol=1e-5), (i, got, want)
    full = batched_ranking_metrics(scores[:, :2], labels[:, :2] * torch.tensor([1.0, 0.0]))
    same = batched_ranking_metrics(scores[:, :2], labels[:, :2] * torch.tensor([1.0, 0.0]), torch.ones(len(rows), 2, dtype=torch.bool))
    assert all(torch.equal(full[k], same[k]) for k in full)
