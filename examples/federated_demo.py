"""All federated modes end to end on planted-signal synthetic data, one client per GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 examples/federated_demo.py

For each mode (gradient averaging with clip + LDP noise, FedAvg with 5 local steps, client/server rounds with a data-less
server rank) the script trains a fresh model for a few hundred rounds through `FederatedRunner` and prints validation
AUC / MRR / nDCG before and after, plus whether every rank ends with identical parameters. It is a functional demo, not
a benchmark: the trunk is a 2-layer random-init DistilBERT so it runs in about a minute.
"""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedrec_b200.config import FedRecConfig  # noqa: E402
from fedrec_b200.data import make_synthetic_mind  # noqa: E402
from fedrec_b200.engine.federated import FederatedRunner  # noqa: E402

ROUNDS = int(os.environ.get("DEMO_ROUNDS", "300"))
os.environ.setdefault("FEDREC_QUIET", "1")      # no per-round echo; the summary below is the output


def run(mode: str, **kw):
    rank = int(os.environ.get("RANK", "0"))
    cfg = FedRecConfig(mode=mode, batch_size=32, title_len=30, bert_layers=2, lr=1e-4, dropout=0.2, snapshot_path="",
                       log_dir="", seed=7, steps_per_epoch=1, **kw)      # count steps, not passes over the shard
    d = make_synthetic_mind(num_news=2000, num_train=512, num_valid=128, title_len=30, seed=22, topics=4,
                            generic_frac=0.0, liked_topics=1, client=rank)
    r = FederatedRunner(cfg, news_index=d.news_index, nid2index=d.nid2index, train_sam=d.train_sam, valid_sam=d.valid_sam)
    with torch.no_grad():       # start unsaturated (sigmoid-before-soft-max loss): small news vectors
        r.model.text_encoder.fc.weight.mul_(0.05)
        r.model.text_encoder.fc.bias.zero_()
    before = r.validate()
    t0 = time.time()
    if mode in ("ddp", "grad_avg"):
        r.cfg.total_epochs = ROUNDS
        r.run_gradient_averaging()
    elif mode == "param_avg":
        r.cfg.total_epochs = ROUNDS // max(cfg.local_steps, 1)
        r.run_parameter_averaging()
    else:
        r.cfg.global_epochs = ROUNDS // max(cfg.total_epochs, 1)
        r.run_client_server()
    if torch.cuda.is_available():
        torch.cuda.synchronize()
    dt = time.time() - t0
    after = r.validate()
    p = r.model.flat.param
    ps = [torch.zeros_like(p) for _ in range(r.W)]
    dist.all_gather(ps, p)
    same = all(torch.equal(ps[0], x) for x in ps)
    if rank == 0:
        fmt = lambda v: "loss %.3f auc %.3f mrr %.3f ndcg@5 %.3f ndcg@10 %.3f" % v if v else "n/a (server rank)"  # noqa: E731
        heap = r.symm.describe() if r.symm else "collective"
        heap += ", optimiser pipelined into the step" if r.pipelined else ""
        print(f"== {mode} {kw} | {r.W} ranks | aggregation: {heap} | {ROUNDS} local steps per client in {dt:.1f} s")
        print(f"   before: {fmt(before)}\n   after:  {fmt(after)}\n   identical parameters on every rank: {same}", flush=True)
    if r.heartbeat:
        r.heartbeat.stop()
    r.logger.close()
    return r


def main():
    srv = 1 if int(os.environ.get("WORLD_SIZE", "1")) > 1 else 0
    runs = [("grad_avg", dict(clip_norm=2.0, noise_multiplier=0.0)),
            # privacy/utility: per-coordinate gradients here are ~1e-4, so even sigma*C = 4e-5 per client is visible and
            # the bench setting (sigma 0.05, C 2 → 0.1 per coordinate) drowns them: DP needs far larger cohorts than 4 clients
            ("grad_avg", dict(clip_norm=2.0, noise_multiplier=2e-5)),
            ("grad_avg", dict(clip_norm=2.0, noise_multiplier=0.05)),
            ("param_avg", dict(local_steps=5)),
            ("client_server", dict(total_epochs=5, server_rank=srv)),
            ("client_server", dict(total_epochs=5, server_rank=srv, news_grad_noise=1e-4))]
    for mode, kw in runs:
        run(mode, **kw)
        dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
