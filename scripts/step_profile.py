"""In-situ per-kernel device time of one training step (warm caches, real clocks) via the PyTorch profiler (CUPTI).

Runs with FEDREC_PDL=0: with programmatic dependent launch a kernel becomes resident while its predecessor drains and its
recorded duration then includes that wait, so the per-kernel times would no longer be exclusive (or add up)."""
import os
import sys

os.environ.setdefault("FEDREC_PDL", "0")

import numpy as np
import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from fedrec_b200.config import FedRecConfig  # noqa: E402
from fedrec_b200.engine import TrainStep  # noqa: E402
from fedrec_b200.models import UserModel  # noqa: E402

DEV = "cuda:0"
N = int(sys.argv[1]) if len(sys.argv) > 1 else 8
# the shipped single-GPU loop (bench.py): FederatedRunner in gradient mode, optimiser pipelined into the step
from fedrec_b200.engine.federated import FederatedRunner  # noqa: E402
cfg = FedRecConfig(mode="grad_avg", batch_size=50, title_len=30, num_news=bench.NUM_NEWS, cuda_graph=False, clip_norm=2.0,
                   noise_multiplier=0.05, synthetic=True, snapshot_path="", log_dir="", num_train=8, num_valid=1)
runner = FederatedRunner(cfg, news_index=bench.synthetic_news_index(np), nid2index={"<unk>": 0}, train_sam=[], valid_sam=[])
assert runner.pipelined
step = runner.step
cand_h, his_h = bench.make_batches(np, 6 + N, 0)
ids = bench.pack_ids(torch, cand_h, his_h).to(DEV)
B, C, L = 50, 5, 50
for i in range(6):
    step.run_packed(ids[i], B, C, L)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for i in range(6, 6 + N):
        step.run_packed(ids[i], B, C, L)
    torch.cuda.synchronize()
rows = [(e.key, e.device_time_total / N, e.count / N) for e in prof.key_averages() if e.device_time_total > 0]
rows.sort(key=lambda r: -r[1])
tot = sum(r[1] for r in rows)
print(f"device time per step: {tot:.1f} us over {sum(r[2] for r in rows):.0f} kernels")
for k, us, n in rows[:28]:
    print(f"{us:9.1f} us {100 * us / tot:5.1f}%  x{n:<4.0f} {k[:110]}")
