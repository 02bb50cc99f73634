"""Full-ranking evaluation with precomputed vectors: the inference path upstream sketches but never wires up
(`UserDataset`, dataset.py:92-109, and `evaluation_split`, evaluation_functions.py:33-46).

    python evaluate.py --snapshot_path snapshot.pt --data_dir UserData        # or --synthetic

Encodes every news item once (frozen trunk + head on the sm_100a kernels when a GPU is present), encodes every
validation user from those vectors, and reports mean AUC / MRR / nDCG@5 / nDCG@10 over the impressions, each ranked
against its full candidate list (all positives and all negatives).
"""
import os
import sys

import numpy as np

from fedrec_b200.config import FedRecConfig
from fedrec_b200.data import load_user_data, make_synthetic_mind
from fedrec_b200.models import UserModel
from fedrec_b200.utils import checkpoint as ckpt
from fedrec_b200.utils.metrics import batched_ranking_metrics, evaluation_split


def ranked_on_device(news_vecs, user_vecs, samples, nid2index, batch: int = 4096) -> np.ndarray:
    """``evaluation_split`` without the per-impression host loop: impressions sorted by candidate count, padded per
    batch to a common width, scored with one batched product and ranked by ``batched_ranking_metrics`` with a validity
    mask. One row per impression that has both classes, in the input order (single-class impressions are skipped, as
    upstream's ``compute_amn`` raises on them)."""
    import torch
    dev = news_vecs.device
    keep = [i for i, s in enumerate(samples) if len(s[1]) > 0 and len(s[2]) > 0]
    keep.sort(key=lambda i: len(samples[i][1]) + len(samples[i][2]))
    out = np.zeros((len(samples), 4))
    for b in range(0, len(keep), batch):
        idx = keep[b:b + batch]
        width = max(len(samples[i][1]) + len(samples[i][2]) for i in idx)
        cand = torch.zeros(len(idx), width, dtype=torch.long)
        lab = torch.zeros(len(idx), width)
        valid = torch.zeros(len(idx), width, dtype=torch.bool)
        for j, i in enumerate(idx):
            poss, negs = list(samples[i][1]), list(samples[i][2])
            n = len(poss) + len(negs)
            cand[j, :n] = torch.tensor([nid2index[x] for x in poss + negs], dtype=torch.long)
            lab[j, : len(poss)] = 1.0
            valid[j, :n] = True
        cand, lab, valid = cand.to(dev), lab.to(dev), valid.to(dev)
        u = user_vecs[torch.tensor(idx, device=dev)]
        scores = torch.bmm(news_vecs[cand], u.unsqueeze(-1)).squeeze(-1)
        m = batched_ranking_metrics(scores, lab, valid)
        out[idx] = torch.stack([m["auc"], m["mrr"], m["ndcg5"], m["ndcg10"]], 1).double().cpu().numpy()
    return out[sorted(keep)]


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    numpy_metrics = "--numpy_metrics" in argv             # the per-impression host loop of upstream, as a cross-check
    argv = [a for a in argv if a != "--numpy_metrics"]
    cfg = FedRecConfig.from_argv(argv, [])
    if cfg.synthetic or not os.path.isdir(cfg.data_dir):
        d = make_synthetic_mind(cfg.num_news, cfg.num_train, cfg.num_valid, cfg.title_len, cfg.max_his_len, seed=cfg.seed)
        nid2index, news_index, valid_sam = d.nid2index, d.news_index, d.valid_sam
    else:
        nid2index, news_index, _, valid_sam = load_user_data(cfg.data_dir)
    model = UserModel(None, news_index, cfg.resolved_device(), config=cfg)
    if cfg.snapshot_path and os.path.exists(cfg.snapshot_path):
        info = ckpt.load_snapshot(cfg.snapshot_path, model)
        print(f"loaded {cfg.snapshot_path} (epochs run: {info['EPOCHS_RUN']})")
    news_vecs = model.encode_all_news()
    user_vecs = model.encode_users(valid_sam, nid2index, news_vecs)
    samples = [(s[0], s[1] if isinstance(s[1], (list, tuple)) else [s[1]], s[2], s[3], s[4]) for s in valid_sam]
    if numpy_metrics:
        res = evaluation_split(news_vecs.cpu().numpy(), user_vecs.cpu().numpy(), samples, nid2index)
    else:
        res = ranked_on_device(news_vecs, user_vecs, samples, nid2index)
    auc, mrr, n5, n10 = res.mean(axis=0) if len(res) else (np.nan,) * 4
    print(f"impressions={len(res)} auc={auc:.4f} mrr={mrr:.4f} ndcg@5={n5:.4f} ndcg@10={n10:.4f}")
    return auc, mrr, n5, n10


if __name__ == "__main__":
    main()
