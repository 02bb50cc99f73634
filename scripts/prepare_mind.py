"""Raw MIND files → client shards in the upstream layout (fedrec_b200/data/mind.py).

  python scripts/prepare_mind.py --news MINDsmall_train/news.tsv MINDsmall_dev/news.tsv \
      --behaviors MINDsmall_train/behaviors.tsv --valid_behaviors MINDsmall_dev/behaviors.tsv \
      --vocab bert-base-uncased/vocab.txt --out "UserData_{rank}" --clients 4 --title_len 30
  torchrun --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 Gradient_Averaging_main.py 3 50 1 --data_dir "UserData_{rank}"
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedrec_b200.data import mind  # noqa: E402


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--news", nargs="+", required=True, help="news.tsv of every split (ids are merged)")
    ap.add_argument("--behaviors", required=True, help="training behaviors.tsv")
    ap.add_argument("--valid_behaviors", default=None, help="validation behaviors.tsv (else --valid_fraction of the training samples)")
    ap.add_argument("--valid_fraction", type=float, default=0.05)
    ap.add_argument("--vocab", required=True, help="BERT uncased vocab.txt")
    ap.add_argument("--out", default="UserData_{rank}", help="output directory; {rank} = client index")
    ap.add_argument("--clients", type=int, default=1)
    ap.add_argument("--title_len", type=int, default=30)
    ap.add_argument("--with_abstract", action="store_true", help="tokenise title + abstract instead of the title alone")
    a = ap.parse_args(argv)
    dirs = mind.prepare(a.news, a.behaviors, a.valid_behaviors, a.vocab, a.out, a.clients, a.title_len, a.with_abstract,
                        a.valid_fraction)
    for d in dirs:
        from fedrec_b200.data import load_user_data
        nid2index, news_index, tr, va = load_user_data(d)
        print(f"{d}: {news_index.shape[0]} news rows x {news_index.shape[2]} tokens, {len(tr)} training and {len(va)} validation samples")
    return 0


if __name__ == "__main__":
    sys.exit(main())
