"""Times fr_gemm_bf16 against torch.matmul (cuBLAS) on the model's GEMM shapes. CUDA events, L2 flushed."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedrec_b200.ops import dispatch as D  # noqa: E402

DEV = "cuda:0"
SHAPES = [  # (name, M, N, K, epi)
    ("bert_qkv", 82500, 2304, 768, "bias"), ("bert_out", 82500, 768, 768, "resid"),
    ("bert_ffn1", 82500, 3072, 768, "gelu"), ("bert_ffn2", 82500, 768, 3072, "resid"),
    ("text_fc1", 82500, 384, 768, "tanh"), ("text_fc", 2750, 400, 768, "bias"),
    ("user_qkv", 2500, 1200, 400, "bias"), ("user_fc1", 2500, 200, 400, "tanh"),
    ("square_8k", 8192, 8192, 8192, "bias"),
]


def timeit(fn, flush, iters=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    except OSError:
        pass
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    rows = []
    for name, M, N, K, epi in SHAPES:
        x = (torch.randn(M, K, device=DEV) * 0.5).bfloat16()
        w = (torch.randn(N, K, device=DEV) / K ** 0.5).bfloat16()
        b = torch.randn(N, device=DEV)
        r = torch.randn(M, N, device=DEV).bfloat16() if epi == "resid" else None
        out = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
        ours_med, ours_min = timeit(lambda: D.linear_bf16(x, w, b, epi, resid=r, out=out), flush)
        variants = {}
        if N >= 256 and M >= 4096:
            for cg in (1, 2):
                for te in (-1, 1):
                    t, _ = timeit(lambda: D.linear_bf16(x, w, b, epi, resid=r, out=out, block_n=256, cta_group=cg, tma_epilogue=te), flush, iters=6)
                    variants[f"cg{cg}_{'tma' if te > 0 else 'direct'}_tflops"] = round(2.0 * M * N * K / t / 1e9, 1)
        wt = w.t()
        lib_med, lib_min = timeit(lambda: torch.matmul(x, wt, out=out), flush)
        fl = 2.0 * M * N * K
        rows.append({"name": name, "M": M, "N": N, "K": K, "epi": epi, "ours_ms": ours_med, "ours_min_ms": ours_min,
                     "ours_tflops": fl / ours_med / 1e9, "cublas_plain_ms": lib_med, "cublas_tflops": fl / lib_med / 1e9,
                     "frac_of_measured_bf16_peak": fl / ours_med / 1e9 / peaks.get("bf16_tflops", 1665.4), **variants})
        print(rows[-1], flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rows, open("gpurun_out/gemm_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()
