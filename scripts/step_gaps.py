"""How long the device sits idle between two replays of the step graph, measured in the real loop.

A profiler cannot answer this: under CUPTI every graph launch costs the host several hundred microseconds, so the device
waits for the host and the trace shows a gap that the unprofiled loop does not have. Instead the captured step carries
one %globaltimer timestamp kernel as its first node and one as its last (FEDREC_STEP_TIMESTAMPS=1, engine/fused.py); the
gap between the end of step i and the start of step i+1 is read from device memory after the loop.

Three loops, the same ones bench.py times: ids already on the device (run_packed), ids pulled from host-mapped pinned
memory by the step's first kernel with the loss pushed back by its last (run_host / loss_of), and the same with a
synchronous read of every loss (lag 0). Usage: python scripts/step_gaps.py [steps]
"""
import os
import sys
from collections import deque

os.environ["FEDREC_STEP_TIMESTAMPS"] = "1"

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from fedrec_b200.config import FedRecConfig  # noqa: E402
from fedrec_b200.engine.federated import FederatedRunner  # noqa: E402

DEV = "cuda:0"
N = int(sys.argv[1]) if len(sys.argv) > 1 else 200
cfg = FedRecConfig(mode="grad_avg", batch_size=50, title_len=30, num_news=bench.NUM_NEWS, cuda_graph=True, clip_norm=2.0,
                   noise_multiplier=0.05, synthetic=True, snapshot_path="", log_dir="", num_train=8, num_valid=1)
runner = FederatedRunner(cfg, news_index=bench.synthetic_news_index(np), nid2index={"<unk>": 0}, train_sam=[], valid_sam=[])
assert runner.pipelined
step, fused = runner.step, runner.step._fused
cand_h, his_h = bench.make_batches(np, 8 + N, 0)
ids_pin = bench.pack_ids(torch, cand_h, his_h).pin_memory()
ids_d, ids_host = ids_pin.to(DEV), ids_pin.numpy()
B, C, L = 50, 5, 50
ring, ctr = fused._ts


def report(name):
    torch.cuda.synchronize()
    n = int(ctr.item())
    assert n % 2 == 0 and n // 2 <= ring.numel() // 2
    t = ring[:n].cpu().numpy().astype(np.int64).reshape(-1, 2)[-N:]
    dur = (t[:, 1] - t[:, 0]) / 1e3
    gap = (t[1:, 0] - t[:-1, 1]) / 1e3
    print(f"{name:<46} step (first to last node) median {np.median(dur):7.1f} us | idle between replays: median "
          f"{np.median(gap):6.1f} us, mean {gap.mean():6.1f} us, p95 {np.percentile(gap, 95):6.1f} us, max {gap.max():7.1f} us")
    ctr.zero_()


for i in range(8):
    step.run_packed(ids_d[i], B, C, L)
torch.cuda.synchronize()
ctr.zero_()
for i in range(8, 8 + N):
    step.run_packed(ids_d[i], B, C, L)
report("ids on the device (run_packed)")

step.enable_host_io(B, C, L, slots=4)
for i in range(3):
    step.loss_of(step.run_host(ids_host[i]))
for lag in (1, 0):
    torch.cuda.synchronize()
    ctr.zero_()
    tickets = deque()
    for i in range(8, 8 + N):
        tickets.append(step.run_host(ids_host[i]))
        if len(tickets) > lag:
            step.loss_of(tickets.popleft())
    while tickets:
        step.loss_of(tickets.popleft())
    report(f"host ids in, loss out every step (lag {lag})")
runner.flush()
