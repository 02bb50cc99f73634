"""Split-K sweep of the weight-gradient GEMM on the step's shapes (L2 flushed between launches).

The step passes the worst-case row count (every id distinct) and the device-side count limits the reduction, so both are
exercised: K_max = 82,500 rows allocated, 890 * 30 = 26,700 effective.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedrec_b200.ops import dispatch as D  # noqa: E402


def main():
    torch.manual_seed(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    shapes = [("text fc1 dW", 82500, 384, 768, 890, 30), ("text fc dW", 2750, 400, 768, 890, 1),
              ("user qkv dW", 2500, 400, 400, None, 1), ("user fc1 dW", 2500, 200, 400, None, 1)]
    for name, K, M1, N1, cnt, scale in shapes:
        A = (torch.randn(K, M1, device="cuda") * 0.1).bfloat16()
        B = (torch.randn(K, N1, device="cuda") * 0.1).bfloat16()
        k_dev = torch.tensor([cnt], device="cuda", dtype=torch.int32) if cnt else None
        if cnt:
            A[cnt * scale:] = 0
        out = torch.zeros(M1, N1, device="cuda")
        for splits in (0, 8, 16, 24, 33, 49, 66):
            D.gemm_grad(A, B, out, k_dev=k_dev, k_scale=scale, splits=splits)
            torch.cuda.synchronize()
            ts = []
            for _ in range(10):
                flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                D.gemm_grad(A, B, out, k_dev=k_dev, k_scale=scale, splits=splits)
                b.record()
                torch.cuda.synchronize()
                ts.append(a.elapsed_time(b) * 1e3)
            ts.sort()
            print(f"{name:12s} K={K} eff={cnt * scale if cnt else K} [{M1}x{N1}] splits={splits:3d} median_us={ts[5]:.1f}", flush=True)


if __name__ == "__main__":
    main()
