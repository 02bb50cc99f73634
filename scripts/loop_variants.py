"""What the per-step loop costs around the captured graph: full loop vs graph + Adam vs graph only vs the input copies alone.

Device time from CUDA events and host enqueue time per step, no profiler attached (scripts/step_gaps.py shows two large
gaps before each graph replay; this script shows they are the profiler's launch overhead, not part of a normal run).
"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from fedrec_b200.config import FedRecConfig
from fedrec_b200.engine import TrainStep
from fedrec_b200.models import UserModel
DEV = "cuda:0"
cfg = FedRecConfig(batch_size=50, title_len=30, num_news=bench.NUM_NEWS, cuda_graph=True, clip_norm=2.0, noise_multiplier=0.05)
model = UserModel(None, bench.synthetic_news_index(np), DEV, config=cfg)
step = TrainStep(model, cfg)
K = 60
cand_h, his_h = bench.make_batches(np, K, 0)
cand = torch.from_numpy(cand_h).to(torch.int32).to(DEV); his = torch.from_numpy(his_h).to(torch.int32).to(DEV)
for i in range(8):
    step.run(cand[i], his[i]); model.adam_step()
torch.cuda.synchronize()
fs = step._fused
g = list(fs._graphs.values())[0]
def timed(fn, n=K):
    torch.cuda.synchronize(); a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); a.record()
    for i in range(n): fn(i)
    b.record(); t1 = time.perf_counter(); torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / n, (t1 - t0) * 1e6 / n
def full(i): step.run(cand[i], his[i]); model.adam_step()
def nocopy(i): g["graph"].replay(); model.adam_step()
def graph_only(i): g["graph"].replay()
def copies_only(i): g["cand"].copy_(cand[i], non_blocking=True); g["his"].copy_(his[i], non_blocking=True)
for name, fn in (("full", full), ("graph+adam", nocopy), ("graph only", graph_only), ("copies only", copies_only), ("full", full)):
    gpu, cpu = timed(fn)
    print(f"{name:12s} gpu {gpu:8.1f} us/step   cpu enqueue {cpu:8.1f} us/step", flush=True)
