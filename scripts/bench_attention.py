"""Isolated timing of the tcgen05 title attention (L2 flushed between launches).

Usage: python scripts/bench_attention.py [n_titles]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedrec_b200.ops import dispatch as D  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 900
    torch.manual_seed(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for T in (30, 50):
        qkv = (torch.randn(n * T, 2304, device="cuda") * 0.5).bfloat16()
        mask = (torch.rand(n, T, device="cuda") < 0.8).int()
        mask[:, 0] = 1
        out = torch.empty(n * T, 768, device="cuda", dtype=torch.bfloat16)
        D.bert_attention(qkv, mask, None, None, n, T, 12, out, impl="tc")
        torch.cuda.synchronize()
        ts = []
        for _ in range(20):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            D.bert_attention(qkv, mask, None, None, n, T, 12, out, impl="tc")
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b) * 1e3)
        ts.sort()
        print(f"bert_attention_tc n_titles={n} T={T} heads=12 median_us={ts[len(ts) // 2]:.1f}", flush=True)


if __name__ == "__main__":
    main()
