"""Aggregation sweep (BASELINE.json config 5): fused NVLink kernels vs NCCL all-reduce, 1 MB – 1 GB.

Launch: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_aggregate.py
Timing: CUDA events on the launching stream, barrier before every iteration, median over iterations, MAX over ranks.
Bus bandwidth uses the usual all-reduce convention 2*(N-1)/N * bytes / time.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedrec_b200 import ops  # noqa: E402
from fedrec_b200.parallel import init_distributed  # noqa: E402
from fedrec_b200.parallel.symm import SymmetricHeap  # noqa: E402

SIZES_MB = [1, 4.66, 16, 64, 270.1, 1024]


def timed(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    t = torch.tensor([ts[len(ts) // 2]], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    info = init_distributed("nccl")
    W, rank = info.world_size, info.rank
    heap = SymmetricHeap.create(info, heap_bytes=int(1100 * 2 ** 20))
    heap.finalize()
    rows = []
    blocks_list = [int(b) for b in os.environ.get("FEDREC_AGG_BLOCKS", "0").split(",")]
    big = heap.alloc_f32(int(1024 * 2 ** 20 / 4))
    for mb in SIZES_MB:
        n = int(mb * 2 ** 20 / 4) // 4 * 4
        buf = big[:n]
        ref = torch.randn(n, device="cuda") * 0.01
        plain = ref.clone()
        bytes_ = n * 4
        busf = 2 * (W - 1) / W * bytes_
        rec = {"mb": mb, "n": n, "world": W}
        # correctness vs NCCL once per size (no clip/noise)
        buf.copy_(ref * (rank + 1))
        chk = (ref * (rank + 1)).clone()
        dist.all_reduce(chk); chk /= W
        heap.grad_allreduce_(buf, clip=0.0, noise_std=0.0, seed=1, scale=1.0 / W)
        torch.cuda.synchronize()
        rec["max_abs_err_vs_nccl"] = float((buf - chk).abs().max())
        for blocks in blocks_list:
            tag = f"_b{blocks}" if blocks else ""
            t = timed(lambda: heap.grad_allreduce_(buf, clip=2.0, noise_std=0.1, seed=3, scale=1.0 / W, blocks=blocks))
            rec["fused_grad_clip_noise_us" + tag] = t
            rec["fused_grad_busbw_gbs" + tag] = busf / t / 1e3
            w = [1.0 / W] * W
            t = timed(lambda: heap.param_average_(buf, w, blocks=blocks))
            rec["fused_param_avg_us" + tag] = t
            rec["fused_param_busbw_gbs" + tag] = busf / t / 1e3
        t = timed(lambda: (dist.all_reduce(plain), plain.mul_(1.0 / W)))
        rec["nccl_allreduce_mean_us"] = t
        rec["nccl_busbw_gbs"] = busf / t / 1e3

        def nccl_clip_noise():
            ops.clip_noise_(plain, 2.0, 0.1, seed=3, stream_id=rank)
            dist.all_reduce(plain)
            plain.mul_(1.0 / W)
        rec["nccl_plus_clip_noise_kernels_us"] = timed(nccl_clip_noise)
        rec["symm_backend"] = heap.backend
        rec["nvls"] = bool(heap.mc_ptr)
        rows.append(rec)
        if rank == 0:
            print(json.dumps(rec), flush=True)
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        tag = "" if heap.backend == "ipc" else ("_nvls" if heap.mc_ptr else "_vmm")
        json.dump(rows, open(f"gpurun_out/agg_sweep_w{W}{tag}.json", "w"), indent=1)
    heap.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
