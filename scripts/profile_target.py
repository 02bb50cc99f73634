"""Tiny single-GPU drivers for `ncu --set full` captures (one kernel family per invocation).

usage: python scripts/profile_target.py gemm_ffn1|gemm_out|gemm_qkv|gemm_ffn2|attention|gemm_grad|layernorm|aggregate|step
       gemm_fold_qkv|gemm_fold_ffn1|gemm_fold_out|gemm_fold_ffn2   (the LayerNorm-folding variants the trunk runs by default)
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedrec_b200.ops import dispatch as D  # noqa: E402

DEV = "cuda:0"
what = sys.argv[1] if len(sys.argv) > 1 else "gemm_ffn1"
M = int(os.environ.get("FEDREC_PROFILE_M", "27000"))
torch.manual_seed(0)

if what.startswith("gemm_fold_"):
    N, K, epi = {"gemm_fold_ffn1": (3072, 768, "gelu"), "gemm_fold_out": (768, 768, "resid"),
                 "gemm_fold_qkv": (2304, 768, "bias"), "gemm_fold_ffn2": (768, 3072, "resid")}[what]
    x = (torch.randn(M, K, device=DEV) * 0.5).bfloat16()
    w = (torch.randn(N, K, device=DEV) / K ** 0.5).bfloat16()
    b = torch.randn(N, device=DEV)
    out = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
    kw = {}
    if epi == "resid":                       # residual normalised on the fly + row statistics of the output
        r = torch.randn(M, N, device=DEV).bfloat16()
        rf = r.float().view(M, N // 64, 64)
        kw = dict(resid=r, r_stats=torch.stack([rf.sum(2), ((rf - rf.mean(2, keepdim=True)) ** 2).sum(2)], 2).contiguous(),
                  r_gamma=torch.rand(N, device=DEV) + 0.5, r_beta=torch.randn(N, device=DEV) * 0.1,
                  stats_out=torch.empty(M, N // 64, 2, device=DEV))
    else:                                    # LayerNorm of the input applied in the epilogue
        xf = x.float().view(M, K // 64, 64)
        kw = dict(a_stats=torch.stack([xf.sum(2), ((xf - xf.mean(2, keepdim=True)) ** 2).sum(2)], 2).contiguous(),
                  col_c=w.float().sum(1).contiguous())
    for _ in range(6):
        D.linear_bf16_fold(x, w, b, epi, out, **kw)
elif what.startswith("gemm_") and what != "gemm_grad":
    N, K, epi = {"gemm_ffn1": (3072, 768, "gelu"), "gemm_out": (768, 768, "resid"), "gemm_qkv": (2304, 768, "bias"),
                 "gemm_ffn2": (768, 3072, "resid")}[what]
    x = (torch.randn(M, K, device=DEV) * 0.5).bfloat16()
    w = (torch.randn(N, K, device=DEV) / K ** 0.5).bfloat16()
    b = torch.randn(N, device=DEV)
    r = torch.randn(M, N, device=DEV).bfloat16() if epi == "resid" else None
    out = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
    for _ in range(6):
        D.linear_bf16(x, w, b, epi, resid=r, out=out)
elif what == "attention":
    n, T, H = M // 30, 30, 12
    qkv = (torch.randn(n * T, 2304, device=DEV) * 0.7).bfloat16()
    out = torch.zeros(n * T, 768, device=DEV, dtype=torch.bfloat16)
    for _ in range(6):
        D.bert_attention(qkv, None, None, None, n, T, H, out)
elif what in ("ue_attention", "ue_attention_bwd"):
    B, L, H, dk = 50, 50, 20, 20
    qkv = (torch.randn(B * L, 3 * H * dk, device=DEV) * 0.8).bfloat16()
    ctx = torch.zeros(B * L, H * dk, device=DEV, dtype=torch.bfloat16)
    stats = torch.zeros(B * H * L, 2, device=DEV)
    dctx = (torch.randn(B * L, H * dk, device=DEV) * 0.1).bfloat16()
    dqkv = torch.zeros_like(qkv)
    for _ in range(6):
        D.ue_mha_fwd(qkv, None, ctx, stats, B, L, H, dk, False, impl="tc")
        D.ue_mha_bwd(qkv, dctx, stats, None, dqkv, B, L, H, dk, False, impl="tc")
elif what == "layernorm":
    x = torch.randn(M, 768, device=DEV).bfloat16()
    g, b = torch.randn(768, device=DEV), torch.randn(768, device=DEV)
    out = torch.empty_like(x)
    for _ in range(6):
        D.layernorm_bf16(x, g, b, out=out)
elif what == "aggregate":
    from fedrec_b200.parallel.symm import SymmetricHeap
    W, n = 8, 1_164_896
    hs = SymmetricHeap.local(W, heap_bytes=8 << 20)
    bufs = [h.alloc_f32(n) for h in hs]
    for b_ in bufs:
        b_.normal_()
    ps = [torch.randn(n, device=DEV) for _ in range(W)]
    ms = [torch.zeros(n, device=DEV) for _ in range(W)]
    vs = [torch.zeros(n, device=DEV) for _ in range(W)]
    for it in range(4):      # one aggregation = sumsq, publish, wait, reduce, wait, adam, commit (7 launches)
        hs[0].grad_allreduce_adam_(bufs[0], clip=2.0, noise_std=0.1, seed=3, scale=1.0 / W, params=ps, exp_avg=ms,
                                   exp_avg_sq=vs, step=it + 1, lr=5e-5, b1=0.9, b2=0.999, eps=1e-8, ranks_here=W)
elif what == "step":
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.data import make_synthetic_mind
    from fedrec_b200.engine import TrainStep
    from fedrec_b200.models import UserModel
    cfg = FedRecConfig(batch_size=50, title_len=30, num_news=8192, cuda_graph=False)
    data = make_synthetic_mind(num_news=8192, num_train=8, num_valid=2, title_len=30, seed=0)
    model = UserModel(None, data.news_index, DEV, config=cfg)
    step = TrainStep(model, cfg)
    for i in range(4):
        cand = torch.randint(1, 8192, (50, 5), device=DEV, dtype=torch.int32)
        his = torch.randint(1, 8192, (50, 50), device=DEV, dtype=torch.int32)
        step.run(cand, his, train=True)
        model.adam_step()
elif what == "gemm_grad":
    A = (torch.randn(M, 384, device=DEV) * 0.3).bfloat16()
    B = (torch.randn(M, 768, device=DEV) * 0.3).bfloat16()
    out = torch.zeros(384, 768, device=DEV)
    for _ in range(6):
        D.gemm_grad(A, B, out)
torch.cuda.synchronize()
print("done", what)
