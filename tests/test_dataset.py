import numpy as np
import torch

from fedrec_b200.data import (BatchSampler, ImpressionTable, NewsDataset, NewsPartDataset, NewsUpdatorDataset,
                              TrainDataset, UserDataset, newsample, make_synthetic_mind)


def test_newsample_pads_with_unk():
    assert newsample(["a", "b"], 4) == ["a", "b", "<unk>", "<unk>"]
    s = newsample(list("abcdefg"), 4)
    assert len(s) == 4 and len(set(s)) == 4


def test_train_dataset_layout(tiny_data):
    ds = TrainDataset(tiny_data.train_sam, tiny_data.nid2index, tiny_data.news_index, seed=0)
    cand, his, label = ds[0]
    _, pos, neg, h, _ = tiny_data.train_sam[0]
    assert cand.shape == (5,) and cand.dtype == np.int64
    assert cand[0] == tiny_data.nid2index[pos]                       # positive first
    assert set(cand[1:]) <= {tiny_data.nid2index[n] for n in neg} | {0}
    assert his.shape == (50,) and label == 0
    k = min(len(h), 50)
    assert list(his[:k]) == [tiny_data.nid2index[n] for n in h][-50:][:k]
    assert (his[k:] == 0).all()


def test_history_is_truncated_to_most_recent():
    d = make_synthetic_mind(num_news=100, num_train=40, num_valid=2, title_len=10, max_his_len=8, seed=1)
    ds = TrainDataset(d.train_sam, d.nid2index, d.news_index, max_his_len=8, seed=0)
    long = [i for i, s in enumerate(d.train_sam) if len(s[3]) > 8]
    assert long, "generator should produce over-long histories (upstream toy shard has 76 > 50)"
    _, his, _ = ds[long[0]]
    assert his.shape == (8,)
    assert list(his) == [d.nid2index[n] for n in d.train_sam[long[0]][3]][-8:]


def test_synthetic_layout(tiny_data):
    ni = tiny_data.news_index
    assert ni.dtype == np.int64 and ni.shape == (300, 2, 12)
    assert (ni[0] == 0).all() and tiny_data.nid2index["<unk>"] == 0
    assert ((ni[:, 1] == 0) | (ni[:, 1] == 1)).all()
    s = tiny_data.train_sam[0]
    assert len(s) == 5 and isinstance(s[1], str) and isinstance(s[2], list) and isinstance(s[3], list)


def test_small_datasets(tiny_data):
    ni = tiny_data.news_index
    assert len(NewsDataset(ni)) == 300 and (NewsDataset(ni)[3] == ni[3]).all()
    nid, rec = NewsPartDataset(ni, [5, 7])[1]
    assert nid == 7 and (rec == ni[7]).all()
    g = torch.randn(2, 400)
    rec, gr = NewsUpdatorDataset(ni, [5, 7], g)[0]
    assert (rec == ni[5]).all() and torch.equal(gr, g[0])
    vecs = np.random.randn(300, 16)
    u = UserDataset(None, tiny_data.valid_sam, vecs, tiny_data.nid2index)
    assert u[0].shape == (50, 16)


def test_batch_sampler(tiny_data):
    tab = ImpressionTable(tiny_data.train_sam, tiny_data.nid2index)
    bs = BatchSampler(tab, 16, seed=0, pin=False)
    seen = 0
    for _ in range(6):
        cand, his = bs.next()
        assert cand.shape == (16, 5) and his.shape == (16, 50) and cand.dtype == torch.int32
        assert (cand >= 0).all() and (cand < 300).all()
        seen += 16
    assert bs.epoch >= 2                                              # 96 draws over a 48-sample shard
    # every sampled negative comes from that impression's negative list (or is the <unk> pad)
    pos2neg = {}
    for s in tiny_data.train_sam:
        pos2neg.setdefault(tiny_data.nid2index[s[1]], set()).update(tiny_data.nid2index[n] for n in s[2])
    for row in cand.numpy():
        assert set(row[1:].tolist()) <= pos2neg[int(row[0])] | {0}


def test_device_prefetcher_yields_the_sampler_sequence(tiny_data):
    from fedrec_b200.data import DevicePrefetcher
    tab = ImpressionTable(tiny_data.train_sam, tiny_data.nid2index)
    ref = BatchSampler(tab, 8, seed=3, pin=False)
    pre = DevicePrefetcher(BatchSampler(tab, 8, seed=3, pin=False), torch.device("cpu"))
    for _ in range(5):
        c0, h0 = ref.next()
        c1, h1 = pre.next()
        assert torch.equal(c0, c1) and torch.equal(h0, h1)


def test_negative_sampling_is_distinct_uniform_and_pads_short_lists():
    """Vectorised Floyd sampling in BatchSampler: every negative equally likely in every slot, no repeats within an
    impression; fewer than npratio negatives → all of them, then <unk> (newsample, dataset.py:10-14)."""
    from collections import Counter
    from fedrec_b200.data import BatchSampler, ImpressionTable
    sam = [(0, "a", [f"n{i}" for i in range(6)], ["a"], "u"), (1, "a", ["n0", "n1"], ["a"], "u"), (2, "a", [], ["a"], "u")]
    n2i = {"<unk>": 0, "a": 1, **{f"n{i}": 2 + i for i in range(6)}}
    s = BatchSampler(ImpressionTable(sam, n2i), 3, seed=1, pin=False)
    any_slot, first_slot, rounds = Counter(), Counter(), 6000
    for _ in range(rounds):
        cand, _ = s.next()
        for row in cand.numpy().tolist():
            assert row[0] == 1
            negs = row[1:]
            if 0 not in negs:
                assert len(set(negs)) == 4 and all(2 <= x <= 7 for x in negs)
                any_slot.update(negs)
                first_slot[negs[0]] += 1
            elif negs.count(0) == 2:
                assert negs[:2] == [2, 3]
            else:
                assert negs == [0, 0, 0, 0]
    assert sum(first_slot.values()) == rounds
    for v in range(2, 8):
        assert abs(any_slot[v] / rounds - 4 / 6) < 0.03 and abs(first_slot[v] / rounds - 1 / 6) < 0.02
