"""Rank every impression of a MIND behaviors file with a trained snapshot and write the leaderboard's prediction format
(`<impression id> [rank of candidate 1, rank of candidate 2, ...]`, 1 = highest score); with labelled files also prints
mean AUC / MRR / nDCG@5 / nDCG@10 (evaluation_functions.py:5-46).

  python scripts/predict_mind.py --behaviors MINDsmall_dev/behaviors.tsv --data_dir UserData_0 \
      --snapshot_path snapshot.pt --out prediction.txt [--bert_path DIR]
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedrec_b200.config import FedRecConfig  # noqa: E402
from fedrec_b200.data import load_user_data, mind  # noqa: E402
from fedrec_b200.engine.serving import Recommender  # noqa: E402
from fedrec_b200.models import UserModel  # noqa: E402
from fedrec_b200.utils import checkpoint as ckpt  # noqa: E402
from fedrec_b200.utils.metrics import compute_amn  # noqa: E402


def predict(behaviors: str, out_path: str, cfg: FedRecConfig, batch: int = 512):
    nid2index, news_index, _, _ = load_user_data(cfg.data_dir)
    model = UserModel(None, news_index, cfg.resolved_device(), config=cfg)
    if cfg.snapshot_path and os.path.exists(cfg.snapshot_path):
        ckpt.load_snapshot(cfg.snapshot_path, model)
    rec = Recommender(model.eval()).refresh()
    imps = mind.read_impressions(behaviors, nid2index)
    metrics = []
    with open(out_path, "w") as f:
        for i in range(0, len(imps), batch):
            chunk = imps[i:i + batch]
            his = rec.pad_histories([[nid2index[n] for n in h] for _, _, h, _, _ in chunk])
            width = max(len(c) for _, _, _, c, _ in chunk)
            cand = torch.zeros(len(chunk), width, dtype=torch.long)
            for j, (_, _, _, c, _) in enumerate(chunk):
                cand[j, : len(c)] = torch.tensor([nid2index.get(n, 0) for n in c], dtype=torch.long)
            scores = rec.score(his, cand).cpu().numpy()
            for j, (imp_id, _, _, c, labels) in enumerate(chunk):
                s = scores[j, : len(c)]
                order = np.argsort(-s, kind="stable")
                ranks = np.empty(len(c), dtype=np.int64)
                ranks[order] = np.arange(1, len(c) + 1)
                f.write(f"{imp_id} [{','.join(map(str, ranks.tolist()))}]\n")
                if labels is not None and 0 < sum(labels) < len(labels):
                    metrics.append(compute_amn(np.asarray(labels), s))
    res = np.asarray(metrics).mean(axis=0) if metrics else None
    return len(imps), res


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    opts = {"--behaviors": None, "--out": "prediction.txt"}
    rest = []
    i = 0
    while i < len(argv):
        if argv[i] in opts and i + 1 < len(argv):
            opts[argv[i]] = argv[i + 1]
            i += 2
        else:
            rest.append(argv[i])
            i += 1
    if not opts["--behaviors"]:
        print(__doc__)
        return 2
    cfg = FedRecConfig.from_argv(rest, [])
    n, res = predict(opts["--behaviors"], opts["--out"], cfg)
    msg = f"{n} impressions ranked -> {opts['--out']}"
    if res is not None:
        msg += f"; auc={res[0]:.4f} mrr={res[1]:.4f} ndcg@5={res[2]:.4f} ndcg@10={res[3]:.4f}"
    print(msg)
    return 0


if __name__ == "__main__":
    sys.exit(main())
