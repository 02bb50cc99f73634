"""Where a CUDA-graph step spends time that is not kernel time: idle gaps on the device timeline.

Profiles a few graph replays with the PyTorch profiler (CUPTI), merges the kernel intervals of all streams and prints the
idle time per step plus the largest gaps with the kernels around them.  Usage: python scripts/step_gaps.py [steps]
"""
import json
import os
import sys
import tempfile

import numpy as np
import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from fedrec_b200.config import FedRecConfig  # noqa: E402
from fedrec_b200.engine import TrainStep  # noqa: E402
from fedrec_b200.models import UserModel  # noqa: E402

DEV = "cuda:0"
N = int(sys.argv[1]) if len(sys.argv) > 1 else 4
# the shipped single-GPU loop (bench.py): FederatedRunner in gradient mode, optimiser pipelined into the step
from fedrec_b200.engine.federated import FederatedRunner  # noqa: E402
cfg = FedRecConfig(mode="grad_avg", batch_size=50, title_len=30, num_news=bench.NUM_NEWS, cuda_graph=True, clip_norm=2.0,
                   noise_multiplier=0.05, synthetic=True, snapshot_path="", log_dir="", num_train=8, num_valid=1)
runner = FederatedRunner(cfg, news_index=bench.synthetic_news_index(np), nid2index={"<unk>": 0}, train_sam=[], valid_sam=[])
assert runner.pipelined
step = runner.step
cand_h, his_h = bench.make_batches(np, 8 + N, 0)
ids = bench.pack_ids(torch, cand_h, his_h).to(DEV)
B, C, L = 50, 5, 50
for i in range(8):
    step.run_packed(ids[i], B, C, L)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(8, 8 + N):
        step.run_packed(ids[i], B, C, L)
    torch.cuda.synchronize()
path = os.path.join(tempfile.mkdtemp(), "trace.json")
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "dur" in e]
ev.sort(key=lambda e: e["ts"])
t0, t1 = ev[0]["ts"], max(e["ts"] + e["dur"] for e in ev)
busy_end, idle, gaps = ev[0]["ts"], 0.0, []
last = ev[0]
for e in ev:
    if e["ts"] > busy_end:
        g = e["ts"] - busy_end
        idle += g
        gaps.append((g, last["name"][:70], e["name"][:70]))
    if e["ts"] + e["dur"] > busy_end:
        busy_end, last = e["ts"] + e["dur"], e
ksum = sum(e["dur"] for e in ev)
print(f"{N} steps: wall {(t1 - t0) / N:.1f} us/step, kernel-time sum {ksum / N:.1f} us/step, "
      f"device idle {idle / N:.1f} us/step in {len(gaps) / N:.0f} gaps/step")
hist = {}
for g, a, b in gaps:
    key = (a.split("(")[0].split("<")[0], b.split("(")[0].split("<")[0])
    n, tot = hist.get(key, (0, 0.0))
    hist[key] = (n + 1, tot + g)
print("idle time by (kernel before -> kernel after), per step:")
for (a, b), (n, tot) in sorted(hist.items(), key=lambda kv: -kv[1][1])[:25]:
    print(f"  {tot / N:7.1f} us  x{n / N:<5.1f} avg {tot / n:5.2f} us   {a}  ->  {b}")
