"""Raw MIND files → upstream shard layout (data/mind.py): tokeniser parity with the `tokenizers` library, the sample
layout of dataset.py:69-86, per-user sharding, and a training step on the result."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedrec_b200.data import ImpressionTable, BatchSampler, TrainDataset, load_user_data, mind  # noqa: E402

VOCAB = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]", "the", "quick", "brown", "fox", "jump", "##s", "##ed", "over", "lazy",
         "dog", ",", ".", "!", "?", "'", "s", "-", "un", "##believ", "##able", "cafe", "naive", "new", "york", "##er", "5",
         "##0", "%", "$", "1", "##00", "a", "b", "##c", "中", "国", "wins", "cup", "rain", "today", "stocks", "fall"]
TEXTS = ["The quick brown fox jumps over the lazy dog's Café, unbelievable! 50% $100 中国 New-Yorker abc xyz",
         "  naïve\tfox jumped... over?!  ", "", "a" * 150 + " fox", "fox\x00\x07 dog � s", "UNBELIEVABLE-fox's"]


@pytest.fixture()
def vocab_file(tmp_path):
    p = tmp_path / "vocab.txt"
    p.write_text("\n".join(VOCAB) + "\n", encoding="utf-8")
    return str(p)


def test_wordpiece_matches_the_tokenizers_library(vocab_file):
    tokenizers = pytest.importorskip("tokenizers")
    ref = tokenizers.BertWordPieceTokenizer(vocab_file, lowercase=True)
    tok = mind.WordPiece.from_file(vocab_file)
    for text in TEXTS:
        assert ["[CLS]"] + tok.tokenize(text) + ["[SEP]"] == ref.encode(text).tokens, text


def test_encode_truncates_pads_and_masks(vocab_file):
    tok = mind.WordPiece.from_file(vocab_file)
    ids, mask = tok.encode("the quick brown fox", 8)
    assert ids.tolist() == [2, 5, 6, 7, 8, 3, 0, 0] and mask.tolist() == [1, 1, 1, 1, 1, 1, 0, 0]
    ids, mask = tok.encode("the quick brown fox jumps over the lazy dog", 6)
    assert ids.tolist() == [2, 5, 6, 7, 8, 3] and mask.tolist() == [1] * 6       # [CLS] + 4 tokens + [SEP]
    ids, mask = tok.encode("", 4)
    assert ids.tolist() == [2, 3, 0, 0] and mask.tolist() == [1, 1, 0, 0]


def _write_mind(tmp_path):
    news = tmp_path / "news.tsv"
    news.write_text("\n".join([
        "N1\tsports\tsoccer\tFox wins cup\tabstract one\turl\t[]\t[]",
        "N2\tweather\t\tRain today\tabstract two\turl\t[]\t[]",
        "N3\tfinance\t\tStocks fall\t\turl\t[]\t[]",
        "N4\tnews\t\tThe lazy dog\t\turl\t[]\t[]",
        "N1\tsports\tsoccer\tduplicate id is ignored\t\turl\t[]\t[]",
    ]) + "\n", encoding="utf-8")
    beh = tmp_path / "behaviors.tsv"
    beh.write_text("\n".join([
        "1\tU1\t11/11/2019 9:05:58 AM\tN1 N2 N9\tN3-1 N4-0 N2-0",
        "2\tU2\t11/11/2019 9:06:58 AM\t\tN1-0 N2-1 N3-1 N9-0",
        "3\tU1\t11/11/2019 9:07:58 AM\tN3\tN4-0 N1-0",
        "4\tU3\t11/11/2019 9:08:58 AM\tN4\tN9-1 N1-0",
    ]) + "\n", encoding="utf-8")
    return str(news), str(beh)


def test_samples_have_the_upstream_layout(tmp_path, vocab_file):
    news_path, beh_path = _write_mind(tmp_path)
    tok = mind.WordPiece.from_file(vocab_file)
    news = mind.read_news([news_path])
    assert list(news) == ["N1", "N2", "N3", "N4"] and news["N1"] == "Fox wins cup"
    nid2index, table = mind.build_news_index(news, tok, 6)
    assert nid2index == {"<unk>": 0, "N1": 1, "N2": 2, "N3": 3, "N4": 4} and table.shape == (5, 2, 6)
    assert table[0].sum() == 0 and table[1, 0].tolist() == [2, VOCAB.index("fox"), VOCAB.index("wins"), VOCAB.index("cup"), 3, 0]
    s = mind.build_samples(beh_path, nid2index)
    # impression 1: one positive; impression 2: two positives, unknown N9 dropped; 3: no click; 4: unknown positive
    assert s == [(0, "N3", ["N4", "N2"], ["N1", "N2"], "U1"), (1, "N2", ["N1"], [], "U2"), (2, "N3", ["N1"], [], "U2")]
    TrainDataset(s, nid2index, table)[0]                       # the reference-named dataset accepts them


def test_sharding_keeps_a_user_on_one_client_and_files_load(tmp_path, vocab_file):
    news_path, beh_path = _write_mind(tmp_path)
    dirs = mind.prepare([news_path], beh_path, None, vocab_file, str(tmp_path / "UserData_{rank}"), num_clients=2,
                        title_len=6, valid_fraction=0.0)
    assert len(dirs) == 2
    seen = {}
    total = 0
    for r, d in enumerate(dirs):
        nid2index, table, tr, va = load_user_data(d)
        assert table.shape == (5, 2, 6) and table.dtype == np.int64 and va == []
        total += len(tr)
        for smp in tr:
            assert seen.setdefault(smp[4], r) == r and mind.client_of(smp[4], 2) == r
    assert total == 3
    with pytest.raises(ValueError):
        mind.prepare([news_path], beh_path, None, vocab_file, str(tmp_path / "flat"), num_clients=2)


def test_command_line_and_a_training_step_on_the_result(tmp_path, vocab_file, tiny_bert, solo_group):
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts"))
    import prepare_mind
    news_path, beh_path = _write_mind(tmp_path)
    out = str(tmp_path / "shard")
    assert prepare_mind.main(["--news", news_path, "--behaviors", beh_path, "--vocab", vocab_file, "--out", out,
                              "--title_len", "8", "--valid_fraction", "0.34"]) == 0
    nid2index, table, tr, va = load_user_data(out)
    assert len(tr) == 2 and len(va) == 1
    tab = ImpressionTable(tr, nid2index)
    cand, his = BatchSampler(tab, batch_size=2, seed=0).next()
    assert cand.shape == (2, 5) and his.shape == (2, 50)
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.engine.federated import FederatedRunner
    cfg = FedRecConfig(mode="grad_avg", data_dir=out, title_len=8, batch_size=2, total_epochs=1, dropout=0.0, device="cpu",
                       backend="gloo", snapshot_path="", log_dir="")
    r = FederatedRunner(cfg, bert_config=tiny_bert)
    loss = r.run_gradient_averaging()
    assert np.isfinite(loss) and r.model.flat.step == 1


def test_prediction_file_in_the_leaderboard_format(tmp_path, vocab_file):
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts"))
    import predict_mind
    news_path, beh_path = _write_mind(tmp_path)
    out = str(tmp_path / "shard")
    mind.prepare([news_path], beh_path, None, vocab_file, out, title_len=8)
    imps = mind.read_impressions(beh_path, load_user_data(out)[0])
    assert [i[0] for i in imps] == ["1", "2", "3", "4"] and imps[0][3] == ["N3", "N4", "N2"] and imps[0][4] == [1, 0, 0]
    assert imps[0][2] == ["N1", "N2"] and imps[3][3] == ["N9", "N1"]           # unknown history id dropped, candidates kept
    unl = tmp_path / "test_behaviors.tsv"
    unl.write_text("9\tU1\tt\tN1\tN3 N4 N2\n", encoding="utf-8")
    assert mind.read_impressions(str(unl))[0][3:] == (["N3", "N4", "N2"], None)
    pred = str(tmp_path / "prediction.txt")
    rc = predict_mind.main(["--behaviors", beh_path, "--out", pred, "--data_dir", out, "--title_len", "8", "--device", "cpu",
                            "--bert_layers", "1", "--bert_dim", "64", "--bert_heads", "4", "--bert_ffn", "128",
                            "--snapshot_path", ""])
    assert rc == 0
    lines = open(pred).read().splitlines()
    assert len(lines) == 4 and lines[0].startswith("1 [") and lines[1].startswith("2 [")
    for line, n in zip(lines, (3, 4, 2, 2)):
        ranks = [int(x) for x in line.split(" ", 1)[1].strip("[]").split(",")]
        assert sorted(ranks) == list(range(1, n + 1))
