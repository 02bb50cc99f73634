"""Top-k recommendations from a snapshot, one JSON object per line in, one out (engine/serving.py).

  echo '{"user": "U1", "history": ["N100", "N7"], "k": 5}' | \
      python scripts/recommend.py --snapshot_path snapshot.pt --data_dir UserData_0 [--bert_path DIR]
  → {"user": "U1", "news": ["N31", ...], "scores": [0.71, ...]}

Requests are answered in batches of up to --batch_size lines; unknown news ids in a history are ignored.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedrec_b200.config import FedRecConfig  # noqa: E402
from fedrec_b200.data import load_user_data, make_synthetic_mind  # noqa: E402
from fedrec_b200.engine.serving import Recommender  # noqa: E402
from fedrec_b200.models import UserModel  # noqa: E402
from fedrec_b200.utils import checkpoint as ckpt  # noqa: E402


def serve(lines, out, cfg: FedRecConfig, default_k: int = 10) -> int:
    if cfg.synthetic or not os.path.isdir(cfg.data_dir):
        d = make_synthetic_mind(cfg.num_news, 8, 2, cfg.title_len, cfg.max_his_len, seed=cfg.seed)
        nid2index, news_index = d.nid2index, d.news_index
    else:
        nid2index, news_index, _, _ = load_user_data(cfg.data_dir)
    index2nid = {i: n for n, i in nid2index.items()}
    model = UserModel(None, news_index, cfg.resolved_device(), config=cfg)
    if cfg.snapshot_path and os.path.exists(cfg.snapshot_path):
        ckpt.load_snapshot(cfg.snapshot_path, model)
    rec = Recommender(model.eval()).refresh()
    served, batch = 0, []

    def flush():
        nonlocal served
        if not batch:
            return
        k = max(int(r.get("k", default_k)) for r in batch)
        his = rec.pad_histories([[nid2index[n] for n in r.get("history", []) if n in nid2index] for r in batch])
        idx, sc = rec.recommend(his, k=k, exclude_seen=True)
        for r, rows, scores in zip(batch, idx.tolist(), sc.tolist()):
            kk = int(r.get("k", default_k))
            out.write(json.dumps({"user": r.get("user"), "news": [index2nid[i] for i in rows[:kk]],
                                  "scores": [round(s, 6) for s in scores[:kk]]}) + "\n")
        out.flush()
        served += len(batch)
        batch.clear()

    for line in lines:
        line = line.strip()
        if not line:
            continue
        batch.append(json.loads(line))
        if len(batch) >= max(cfg.batch_size, 1):
            flush()
    flush()
    return served


def main(argv=None) -> int:
    cfg = FedRecConfig.from_argv(list(sys.argv[1:] if argv is None else argv), [])
    serve(sys.stdin, sys.stdout, cfg)
    return 0


if __name__ == "__main__":
    sys.exit(main())
