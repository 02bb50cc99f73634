"""From raw MIND files to the client shards the trainers read.

Upstream ships four pre-built files per client (``bert_news_index.npy``, ``bert_nid2index.pkl``,
``train_sam_uid.pkl``, ``valid_sam_uid.pkl``; main.py:146-157, client.py:214-225) and no code that makes them. This
module is that missing step, so a run can start from the public dataset:

* ``WordPiece``: BERT's uncased tokenisation (clean-up, lower-casing, accent stripping, punctuation and CJK splitting,
  greedy longest-match word pieces) from a ``vocab.txt``; no tokenizer library needed on the training box;
* ``build_news_index``: ``news.tsv`` → ``nid2index`` (row 0 = ``<unk>``) and the ``[N, 2, T]`` int64 array of token ids
  and attention masks the text encoder gathers from (encoder.py:24-28 expects exactly this layout);
* ``build_samples``: ``behaviors.tsv`` → one sample per clicked candidate,
  ``(impression, positive, [negatives], [history], user)``, the item layout of dataset.py:69-86;
* ``shard_samples`` / ``write_shard``: users are dealt to clients by a stable hash, so a user's impressions stay on one
  client (the federated setting, README.md:22), and each client directory is written in the upstream file layout.

``scripts/prepare_mind.py`` is the command line around it.
"""
from __future__ import annotations

import os
import pickle
import unicodedata
import zlib
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

UNK_NEWS = "<unk>"
Sample = Tuple[int, str, List[str], List[str], str]


# ---- tokenisation ------------------------------------------------------------------------------------------------
def _is_whitespace(ch: str) -> bool:
    return ch in " \t\n\r" or unicodedata.category(ch) == "Zs"


def _is_control(ch: str) -> bool:
    if ch in "\t\n\r":
        return False
    return unicodedata.category(ch).startswith("C")


def _is_punctuation(ch: str) -> bool:
    cp = ord(ch)
    if 33 <= cp <= 47 or 58 <= cp <= 64 or 91 <= cp <= 96 or 123 <= cp <= 126:
        return True
    return unicodedata.category(ch).startswith("P")


def _is_cjk(cp: int) -> bool:
    return (0x4E00 <= cp <= 0x9FFF or 0x3400 <= cp <= 0x4DBF or 0x20000 <= cp <= 0x2A6DF or 0x2A700 <= cp <= 0x2B73F
            or 0x2B740 <= cp <= 0x2B81F or 0x2B820 <= cp <= 0x2CEAF or 0xF900 <= cp <= 0xFAFF or 0x2F800 <= cp <= 0x2FA1F)


class WordPiece:
    """BERT-uncased tokenisation from a vocabulary file (one token per line, line number = id)."""

    def __init__(self, vocab: Dict[str, int], lower: bool = True, max_chars_per_word: int = 100):
        self.vocab = vocab
        self.lower = lower
        self.max_chars = max_chars_per_word
        for tok in ("[PAD]", "[UNK]", "[CLS]", "[SEP]"):
            if tok not in vocab:
                raise ValueError(f"vocabulary lacks {tok}")
        self.pad_id, self.unk_id = vocab["[PAD]"], vocab["[UNK]"]
        self.cls_id, self.sep_id = vocab["[CLS]"], vocab["[SEP]"]

    @classmethod
    def from_file(cls, path: str, lower: bool = True) -> "WordPiece":
        vocab: Dict[str, int] = {}
        with open(path, encoding="utf-8") as f:
            for i, line in enumerate(f):
                vocab[line.rstrip("\n")] = i
        return cls(vocab, lower)

    def basic(self, text: str) -> List[str]:
        out = []
        for ch in text:
            cp = ord(ch)
            if cp == 0 or cp == 0xFFFD or _is_control(ch):
                continue
            if _is_cjk(cp):
                out.append(f" {ch} ")
            else:
                out.append(" " if _is_whitespace(ch) else ch)
        words: List[str] = []
        for word in "".join(out).split():
            if self.lower:
                word = "".join(c for c in unicodedata.normalize("NFD", word.lower()) if unicodedata.category(c) != "Mn")
            cur: List[str] = []
            for ch in word:
                if _is_punctuation(ch):
                    if cur:
                        words.append("".join(cur))
                        cur = []
                    words.append(ch)
                else:
                    cur.append(ch)
            if cur:
                words.append("".join(cur))
        return words

    def pieces(self, word: str) -> List[str]:
        if len(word) > self.max_chars:
            return ["[UNK]"]
        out, start = [], 0
        while start < len(word):
            end, found = len(word), None
            while start < end:
                sub = word[start:end] if start == 0 else "##" + word[start:end]
                if sub in self.vocab:
                    found = sub
                    break
                end -= 1
            if found is None:
                return ["[UNK]"]
            out.append(found)
            start = end
        return out

    def tokenize(self, text: str) -> List[str]:
        return [p for w in self.basic(text) for p in self.pieces(w)]

    def encode(self, text: str, length: int) -> Tuple[np.ndarray, np.ndarray]:
        """``[CLS] tokens [SEP]`` truncated and padded to ``length``: (ids, mask), both int64."""
        ids = [self.cls_id] + [self.vocab[t] for t in self.tokenize(text)][: max(length - 2, 0)] + [self.sep_id]
        ids = ids[:length]
        row = np.full(length, self.pad_id, dtype=np.int64)
        row[: len(ids)] = ids
        mask = np.zeros(length, dtype=np.int64)
        mask[: len(ids)] = 1
        return row, mask


# ---- news table ---------------------------------------------------------------------------------------------------
def read_news(paths: Sequence[str], text_columns: Sequence[int] = (3,)) -> Dict[str, str]:
    """``news.tsv`` (id, category, subcategory, title, abstract, url, entities...) of one or more splits → id → text.
    The first occurrence of an id wins, so train and dev files can be given together."""
    out: Dict[str, str] = {}
    for path in paths:
        with open(path, encoding="utf-8") as f:
            for line in f:
                cols = line.rstrip("\n").split("\t")
                if not cols or not cols[0] or cols[0] in out:
                    continue
                out[cols[0]] = " ".join(cols[c] for c in text_columns if c < len(cols) and cols[c])
    return out


def build_news_index(news: Dict[str, str], tok: WordPiece, title_len: int = 30) -> Tuple[Dict[str, int], np.ndarray]:
    """id → row and the ``[N + 1, 2, title_len]`` table; row 0 is the all-padding ``<unk>`` news every unknown id and
    every history pad maps to."""
    nid2index: Dict[str, int] = {UNK_NEWS: 0}
    table = np.zeros((len(news) + 1, 2, title_len), dtype=np.int64)
    for nid, text in news.items():
        i = nid2index[nid] = len(nid2index)
        table[i, 0], table[i, 1] = tok.encode(text, title_len)
    return nid2index, table


# ---- impressions ----------------------------------------------------------------------------------------------------
def build_samples(behaviors_path: str, nid2index: Optional[Dict[str, int]] = None, min_negatives: int = 0,
                  start_index: int = 0) -> List[Sample]:
    """``behaviors.tsv`` (impression, user, time, history, "nid-label" ...) → one sample per clicked candidate with the
    impression's non-clicked candidates as its negatives. Ids missing from ``nid2index`` are dropped from histories and
    negatives; an impression whose positive is unknown is skipped."""
    known = (lambda n: True) if nid2index is None else nid2index.__contains__
    out: List[Sample] = []
    with open(behaviors_path, encoding="utf-8") as f:
        for line in f:
            cols = line.rstrip("\n").split("\t")
            if len(cols) < 5:
                continue
            user, history = cols[1], [n for n in cols[3].split() if known(n)]
            pos, neg = [], []
            for item in cols[4].split():
                nid, _, label = item.rpartition("-")
                if not nid or not known(nid):
                    continue
                (pos if label == "1" else neg).append(nid)
            if len(neg) < min_negatives:
                continue
            for p in pos:
                out.append((start_index + len(out), p, list(neg), list(history), user))
    return out


def read_impressions(behaviors_path: str, nid2index: Optional[Dict[str, int]] = None):
    """``behaviors.tsv`` kept as impressions (for ranking a released split): a list of
    ``(impression_id, user, history, candidates, labels or None)``; labels are None when the file has none (MIND test).
    Unknown history ids are dropped; unknown candidates stay (they rank as the ``<unk>`` news) so that every impression
    keeps its candidate count."""
    known = (lambda n: True) if nid2index is None else nid2index.__contains__
    out = []
    with open(behaviors_path, encoding="utf-8") as f:
        for line in f:
            cols = line.rstrip("\n").split("\t")
            if len(cols) < 5:
                continue
            cands, labels = [], []
            for item in cols[4].split():
                nid, sep, label = item.rpartition("-")
                if sep and label in ("0", "1"):
                    cands.append(nid)
                    labels.append(int(label))
                else:
                    cands.append(item)
            out.append((cols[0], cols[1], [n for n in cols[3].split() if known(n)], cands,
                        labels if len(labels) == len(cands) and cands else None))
    return out


def client_of(user: str, num_clients: int) -> int:
    """Stable across processes and Python versions (unlike ``hash``)."""
    return zlib.crc32(user.encode("utf-8")) % max(num_clients, 1)


def shard_samples(samples: Iterable[Sample], num_clients: int) -> List[List[Sample]]:
    shards: List[List[Sample]] = [[] for _ in range(max(num_clients, 1))]
    for s in samples:
        shards[client_of(s[4], num_clients)].append(s)
    return shards


def write_shard(out_dir: str, nid2index: Dict[str, int], news_index: np.ndarray, train: Sequence[Sample],
                valid: Sequence[Sample]) -> None:
    """The four files of one client in the upstream layout (read back by ``load_user_data``)."""
    os.makedirs(out_dir, exist_ok=True)
    np.save(os.path.join(out_dir, "bert_news_index.npy"), news_index)
    for name, obj in (("bert_nid2index.pkl", nid2index), ("train_sam_uid.pkl", list(train)), ("valid_sam_uid.pkl", list(valid))):
        tmp = os.path.join(out_dir, name + ".tmp")
        with open(tmp, "wb") as f:
            pickle.dump(obj, f, protocol=4)
        os.replace(tmp, os.path.join(out_dir, name))


def prepare(news_paths: Sequence[str], train_behaviors: str, valid_behaviors: Optional[str], vocab_path: str,
            out_pattern: str, num_clients: int = 1, title_len: int = 30, with_abstract: bool = False,
            valid_fraction: float = 0.0) -> List[str]:
    """The whole pipeline. ``out_pattern`` may contain ``{rank}`` (one directory per client, what ``--data_dir
    UserData_{rank}`` expects). Without a validation file the last ``valid_fraction`` of each client's samples is held out.
    Returns the directories written."""
    tok = WordPiece.from_file(vocab_path)
    news = read_news(news_paths, (3, 4) if with_abstract else (3,))
    nid2index, table = build_news_index(news, tok, title_len)
    train = build_samples(train_behaviors, nid2index)
    valid = build_samples(valid_behaviors, nid2index) if valid_behaviors else []
    if num_clients > 1 and "{rank}" not in out_pattern:
        raise ValueError("several clients need {rank} in the output pattern")
    tr_shards, va_shards = shard_samples(train, num_clients), shard_samples(valid, num_clients)
    dirs = []
    for r in range(max(num_clients, 1)):
        tr, va = tr_shards[r], va_shards[r]
        if not valid_behaviors and valid_fraction > 0 and tr:
            cut = max(1, int(round(len(tr) * (1.0 - valid_fraction))))
            tr, va = tr[:cut], tr[cut:]
        d = out_pattern.format(rank=r)
        write_shard(d, nid2index, table, tr, va)
        dirs.append(d)
    return dirs
