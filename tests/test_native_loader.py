"""csrc/host_loader.cu through data/native_loader.py: the C++ batch producer follows the sampling rules of
dataset.py:69-86 / :10-14 (clicked news first, distinct uniformly sampled negatives, <unk> padding, 0-padded most recent
history) and visits every impression once per pass."""
import os
import sys
import time
from collections import Counter

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedrec_b200.data import BatchSampler, ImpressionTable  # noqa: E402
from fedrec_b200.data.native_loader import NativeBatchLoader  # noqa: E402


def _table():
    sam = [(0, "a", [f"n{i}" for i in range(6)], ["a", "n1"], "u"), (1, "n0", ["n0", "n1"], ["a"], "u"),
           (2, "n2", [], [], "u"), (3, "n5", [f"n{i}" for i in range(4)], ["n3"] * 70, "u")]
    n2i = {"<unk>": 0, "a": 1, **{f"n{i}": 2 + i for i in range(6)}}
    return ImpressionTable(sam, n2i), n2i


def test_layout_sampling_rules_and_epoch_coverage():
    tab, _ = _table()
    ld = NativeBatchLoader(tab, batch_size=2, seed=7, depth=3)
    assert ld.n == 2 * (50 + 5)
    pos_seen, any_slot, first_slot, rounds = Counter(), Counter(), Counter(), 6000
    for r in range(rounds):
        cand, his = ld.next()
        assert cand.shape == (2, 5) and his.shape == (2, 50) and cand.dtype.is_floating_point is False
        if r % 2 == 1:
            assert ld.epoch == r // 2 + 1                       # 4 impressions, 2 per batch: a pass every two batches
        for c, h in zip(cand.tolist(), his.tolist()):
            pos_seen[c[0]] += 1
            negs = c[1:]
            if c[0] == 1:                                        # 6 negatives: 4 distinct, uniform
                assert len(set(negs)) == 4 and all(2 <= x <= 7 for x in negs) and h[:2] == [1, 3] and sum(h[2:]) == 0
                any_slot.update(negs)
                first_slot[negs[0]] += 1
            elif c[0] == 2:                                      # 2 negatives: both, then <unk>
                assert negs == [2, 3, 0, 0] and h[0] == 1
            elif c[0] == 4:                                      # none
                assert negs == [0, 0, 0, 0] and sum(h) == 0
            else:                                                # exactly npratio negatives: a permutation of them
                assert c[0] == 7 and sorted(negs) == [2, 3, 4, 5] and h == [5] * 50
    assert set(pos_seen.values()) == {rounds // 2}              # every impression exactly once per pass
    n = rounds // 2
    for v in range(2, 8):
        assert abs(any_slot[v] / n - 4 / 6) < 0.03 and abs(first_slot[v] / n - 1 / 6) < 0.025
    ld.close()


def test_seeded_and_packed_output_matches_the_step_layout():
    tab, _ = _table()
    a, b, c = NativeBatchLoader(tab, 3, seed=1), NativeBatchLoader(tab, 3, seed=1), NativeBatchLoader(tab, 3, seed=2)
    xs = [a.next_packed().copy() for _ in range(20)]
    ys = [b.next_packed().copy() for _ in range(20)]
    zs = [c.next_packed().copy() for _ in range(20)]
    assert all(np.array_equal(x, y) for x, y in zip(xs, ys)) and not all(np.array_equal(x, z) for x, z in zip(xs, zs))
    out = np.zeros(a.n + 5, dtype=np.int32)
    got = a.next_packed(out)
    assert got is out and out[a.n:].sum() == 0
    his, cand = out[: 3 * 50].reshape(3, 50), out[3 * 50: a.n].reshape(3, 5)
    assert set(cand[:, 0].tolist()) <= {1, 2, 4, 7} and his.shape == (3, 50)
    for x in (a, b, c):
        x.close()


def test_bad_arguments_are_refused_and_the_thread_stops():
    tab, _ = _table()
    with pytest.raises(RuntimeError, match="fr_loader_create"):
        NativeBatchLoader(tab, batch_size=2, npratio=100)
    ld = NativeBatchLoader(tab, 2)
    ld.next()
    t0 = time.time()
    ld.close()
    ld.close()
    assert time.time() - t0 < 2.0


def test_it_is_faster_than_the_numpy_sampler(tiny_data):
    tab = ImpressionTable(tiny_data.train_sam, tiny_data.nid2index)
    ld, ref = NativeBatchLoader(tab, 50, seed=0, depth=16), BatchSampler(tab, 50, seed=0, pin=False)
    for _ in range(20):
        ld.next_packed(); ref.next()
    t0 = time.perf_counter()
    for _ in range(300):
        ld.next_packed()
    t_native = time.perf_counter() - t0
    t0 = time.perf_counter()
    for _ in range(300):
        ref.next()
    t_numpy = time.perf_counter() - t0
    ld.close()
    assert t_native < t_numpy, (t_native, t_numpy)


def test_the_runner_trains_from_the_native_loader(tiny_bert, solo_group):
    import torch
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.engine.federated import FederatedRunner
    cfg = FedRecConfig(mode="grad_avg", synthetic=True, num_news=120, num_train=10, num_valid=2, title_len=8, batch_size=4,
                       total_epochs=2, dropout=0.0, device="cpu", backend="gloo", snapshot_path="", log_dir="",
                       native_loader=True)
    r = FederatedRunner(cfg, bert_config=tiny_bert)
    assert r._native is not None
    before = r.model.flat.param.clone()
    loss = r.run_gradient_averaging()
    assert np.isfinite(loss) and r.model.flat.step == 2 * 3 and not torch.equal(before, r.model.flat.param)
    assert r._native.epoch >= 2
    r.close()
    assert r._native is None


@pytest.mark.gpu
def test_on_the_gpu_the_runner_steps_through_the_host_ring(solo_group, monkeypatch):
    """native_loader on the sm_100a path: C++ producer → host-mapped ring → captured step with the pipelined optimiser;
    same number of updates and a loss in the same place as the sampler + prefetch-stream path."""
    import torch
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.engine.federated import FederatedRunner
    from fedrec_b200.models.distilbert import DistilBertConfigLite
    monkeypatch.setenv("FEDREC_QUIET", "1")
    out = {}
    for native in (False, True):
        torch.manual_seed(0)
        cfg = FedRecConfig(mode="grad_avg", synthetic=True, num_news=600, num_train=64, num_valid=4, title_len=30,
                           batch_size=8, total_epochs=2, dropout=0.0, clip_norm=2.0, snapshot_path="", log_dir="",
                           native_loader=native, seed=3)
        r = FederatedRunner(cfg, bert_config=DistilBertConfigLite(n_layers=2))
        assert r.step.impl == "fused" and (r._native is not None) == native
        loss = r.run_gradient_averaging()
        out[native] = (loss, int(r.model.flat.step), r.model.flat.param.clone())
        r.close()
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()
    assert out[True][1] == out[False][1] == 2 * 8
    assert all(l == l and 1.0 < l < 2.2 for l in (out[True][0], out[False][0])), (out[True][0], out[False][0])
    assert not torch.equal(out[True][2], out[False][2])                 # different batch order, both trained
