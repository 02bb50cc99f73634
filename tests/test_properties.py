"""Property tests (hypothesis) for the CPU oracles: metrics invariances, de-dup / segment-sum identities."""
import numpy as np
import torch
from hypothesis import given, settings, strategies as st

from fedrec_b200.ops import reference as ref
from fedrec_b200.utils import metrics as M

# scores on a grid (divided by 10000 in the test): no ties, and none created by an affine map
labels_scores = st.integers(2, 30).flatmap(
    lambda n: st.tuples(st.lists(st.integers(0, 1), min_size=n, max_size=n),
                        st.lists(st.integers(0, 10000), min_size=n, max_size=n, unique=True)))


@settings(max_examples=60, deadline=None)
@given(labels_scores)
def test_ranking_metrics_are_bounded_and_rank_invariant(ls):
    y, s = np.asarray(ls[0]), np.asarray(ls[1], dtype=np.float64) / 10000.0
    if y.sum() in (0, len(y)):
        y[0], y[1] = 1, 0
    a = M.compute_amn(y, s)
    assert all(0.0 <= v <= 1.0 + 1e-12 for v in a)
    b = M.compute_amn(y, 3.0 * s + 1.0)               # metrics depend on the ranking only
    assert np.allclose(a, b, atol=1e-12)
    order = np.argsort(-s)
    if y[order][: y.sum()].all():                      # all positives ranked first: every metric is perfect
        assert np.allclose(a, (1.0, M.mrr_score(y, s), 1.0, 1.0), atol=1e-12)
        assert a[0] == 1.0


@settings(max_examples=60, deadline=None)
@given(st.lists(st.integers(0, 40), min_size=1, max_size=200), st.integers(1, 6))
def test_dedup_and_segment_sum_identities(ids, D):
    t = torch.tensor(ids, dtype=torch.int64)
    uniq, inv = ref.dedup(t)
    assert torch.equal(uniq[inv], t)                   # uniq[inv[k]] == ids[k]
    assert torch.equal(uniq, torch.unique(t))          # sorted, no duplicates
    vals = torch.arange(len(ids) * D, dtype=torch.float32).view(len(ids), D)
    out = ref.segment_sum(vals, inv, uniq.numel())
    assert torch.allclose(out.sum(0), vals.sum(0))     # nothing lost, nothing counted twice
    k = int(inv[0])
    assert torch.allclose(out[k], vals[inv == k].sum(0))
