"""What the fused epilogues cost: fr_gemm_bf16 on the trunk shapes at the in-step row count, per epilogue, against plain
cuBLAS (torch.matmul, no epilogue). CUDA events, L2 flushed between iterations."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedrec_b200.ops import dispatch as D  # noqa: E402
from scripts.bench_gemm import timeit  # noqa: E402

DEV = "cuda:0"
M = int(os.environ.get("M", "26880"))
SHAPES = [("qkv", 2304, 768, ["bias"]), ("out", 768, 768, ["bias", "resid"]), ("ffn1", 3072, 768, ["bias", "gelu", "tanh"]),
          ("ffn2", 768, 3072, ["bias", "resid"]), ("text_fc1", 384, 768, ["bias", "tanh"])]


def main():
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    for name, N, K, epis in SHAPES:
        x = (torch.randn(M, K, device=DEV) * 0.5).bfloat16()
        w = (torch.randn(N, K, device=DEV) / K ** 0.5).bfloat16()
        b = torch.randn(N, device=DEV)
        r = torch.randn(M, N, device=DEV).bfloat16()
        out = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
        fl = 2.0 * M * N * K
        row = {"name": name, "M": M, "N": N, "K": K}
        for epi in epis:
            t, _ = timeit(lambda: D.linear_bf16(x, w, b, epi, resid=r if epi == "resid" else None, out=out), flush)
            row[f"ours_{epi}_us"] = round(t * 1e3, 1)
            row[f"ours_{epi}_tflops"] = round(fl / t / 1e9)
        wt = w.t()
        t, _ = timeit(lambda: torch.matmul(x, wt, out=out), flush)
        row["cublas_us"] = round(t * 1e3, 1)
        row["cublas_tflops"] = round(fl / t / 1e9)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
