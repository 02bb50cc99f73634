"""LayerNorm folded into the GEMM epilogues (fr_gemm_bf16_fold, cfg.ln_fold) against plain fp32 PyTorch."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from fedrec_b200 import ops
from fedrec_b200.ops import dispatch as D

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _stats(x):
    """[M, W/64, 2]: (sum, M2 = sum of squared deviations from the slice mean) of every 64-column slice of every row."""
    xf = x.double().view(x.shape[0], -1, 64)
    return torch.stack([xf.sum(2), ((xf - xf.mean(2, keepdim=True)) ** 2).sum(2)], 2).float().contiguous()


@pytest.mark.parametrize("M,epi", [(4500, "bias"), (4500, "gelu"), (300, "gelu")])
def test_input_layernorm_folded_into_the_gemm(M, epi):
    """out = epi(LN(x)·W0^T + b0) computed from the un-normalised x, its row statistics and gamma-scaled weights."""
    ops.require_native()
    g = torch.Generator(device=DEV).manual_seed(0)
    K, N = 768, 1024
    x = (torch.randn(M, K, device=DEV, generator=g) * 1.5 + 0.7).bfloat16()
    w0 = torch.randn(N, K, device=DEV, generator=g) * 0.03
    b0 = torch.randn(N, device=DEV, generator=g) * 0.1
    gamma = 1.0 + 0.2 * torch.randn(K, device=DEV, generator=g)
    beta = 0.1 * torch.randn(K, device=DEV, generator=g)
    wf = (w0 * gamma[None, :]).bfloat16()
    c = wf.float().sum(1).contiguous()
    bf = (b0 + w0 @ beta).contiguous()
    out = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
    D.linear_bf16_fold(x, wf, bf, epi, out, a_stats=_stats(x), col_c=c, ln_eps=1e-12)
    xf = x.float()
    mu, var = xf.mean(1, keepdim=True), xf.var(1, unbiased=False, keepdim=True)
    same_inputs = ((xf - mu) * torch.rsqrt(var + 1e-12)) @ wf.float().t() + bf      # the kernel's own rounded operands
    true_ln = F.layer_norm(xf, (K,), gamma, beta, 1e-12) @ w0.t() + b0            # what the two-kernel path computes
    if epi == "gelu":
        same_inputs, true_ln = F.gelu(same_inputs), F.gelu(true_ln)
    err = (out.float() - same_inputs).abs()
    assert float(err.max()) < 0.05 and float(err.mean()) < 4e-3, (float(err.max()), float(err.mean()))
    assert float((out.float() - true_ln).abs().mean()) < 8e-3


@pytest.mark.parametrize("M,K,with_ln", [(4500, 768, True), (5000, 3072, True), (4500, 768, False), (200, 768, True)])
def test_residual_layernorm_on_the_fly_and_row_statistics(M, K, with_ln):
    ops.require_native()
    g = torch.Generator(device=DEV).manual_seed(1)
    N = 768
    a = (torch.randn(M, K, device=DEV, generator=g) * 0.5).bfloat16()
    w = (torch.randn(N, K, device=DEV, generator=g) * 0.03).bfloat16()
    b = torch.randn(N, device=DEV, generator=g) * 0.1
    r = (torch.randn(M, N, device=DEV, generator=g) * 2.0 - 0.4).bfloat16()
    gamma = 1.0 + 0.2 * torch.randn(N, device=DEV, generator=g)
    beta = 0.1 * torch.randn(N, device=DEV, generator=g)
    out = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
    st = torch.full((M, N // 64, 2), 123.0, device=DEV)                 # stored, not accumulated: no zeroing needed
    kw = dict(r_stats=_stats(r), r_gamma=gamma, r_beta=beta) if with_ln else {}
    D.linear_bf16_fold(a, w, b, "resid", out, resid=r, stats_out=st, **kw)
    res = F.layer_norm(r.float(), (N,), gamma, beta, 1e-12) if with_ln else r.float()
    want = a.float() @ w.float().t() + b + res
    err = (out.float() - want).abs()
    assert float(err.max()) < 0.06 and float(err.mean()) < 6e-3, (float(err.max()), float(err.mean()))
    torch.testing.assert_close(st, _stats(want), atol=0.08, rtol=3e-3)
    st2 = torch.zeros_like(st)
    D.linear_bf16_fold(a, w, b, "resid", out, resid=r, stats_out=st2, **kw)
    assert torch.equal(st, st2)                                         # no atomics: bit-reproducible


@pytest.mark.parametrize("case", ["offset50", "outlier_channels", "wide_gamma"])
def test_folding_survives_pretrained_like_residual_streams(case):
    """Rows as a pretrained DistilBERT produces them: a large common offset (mean 50, std 1), a few channels 100x larger
    than the rest, LayerNorm gains spanning 0.1..10. E[x^2] - mean^2 in fp32 loses these to cancellation; the (sum, M2)
    slices merged with the parallel-variance formula must not. Checked on the statistics a producer GEMM writes and on
    both consumers (input LayerNorm folded into the GEMM, residual LayerNorm on the fly)."""
    ops.require_native()
    g = torch.Generator(device=DEV).manual_seed(7)
    M, K, N = 2048, 768, 768
    base = torch.randn(M, K, device=DEV, generator=g)
    gamma = 1.0 + 0.2 * torch.randn(K, device=DEV, generator=g)
    if case == "offset50":
        x32 = base + 50.0
    elif case == "outlier_channels":
        x32 = base.clone()
        x32[:, [5, 300, 301, 767]] *= 100.0
    else:
        x32 = base * 1.5 + 0.7
        gamma = torch.exp(torch.empty(K, device=DEV).uniform_(-2.3, 2.3, generator=g))     # 0.1 .. 10
    beta = 0.1 * torch.randn(K, device=DEV, generator=g)
    # (1) a producer writes the row: residual GEMM with a zero product, so out = r exactly and stats_out describes r
    r = x32.bfloat16()
    a0 = torch.zeros(M, 64, device=DEV, dtype=torch.bfloat16)
    w0 = torch.zeros(N, 64, device=DEV, dtype=torch.bfloat16)
    out = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
    st = torch.empty(M, N // 64, 2, device=DEV)
    D.linear_bf16_fold(a0, w0, None, "resid", out, resid=r, stats_out=st)
    assert torch.equal(out, r)
    rf = r.double()
    mu_true, var_true = rf.mean(1), rf.var(1, unbiased=False)
    n = float(K)
    mean_s = st[..., 0].double() / 64.0
    mu = st[..., 0].double().sum(1) / n
    var = (st[..., 1].double() + 64.0 * (mean_s - mu[:, None]) ** 2).sum(1) / n
    assert float(((mu - mu_true).abs() / (mu_true.abs() + 1.0)).max()) < 1e-5
    assert float(((var - var_true).abs() / var_true).max()) < 2e-3, float(((var - var_true).abs() / var_true).max())
    # (2) consumer: LN(x) folded into the GEMM, from the kernel-written statistics
    wlin = torch.randn(512, K, device=DEV, generator=g) * 0.03
    blin = torch.randn(512, device=DEV, generator=g) * 0.1
    wf = (wlin * gamma[None, :]).bfloat16()
    c = wf.float().sum(1).contiguous()
    bfold = (blin + wlin @ beta).contiguous()
    o2 = torch.empty(M, 512, device=DEV, dtype=torch.bfloat16)
    D.linear_bf16_fold(r, wf, bfold, "bias", o2, a_stats=st, col_c=c, ln_eps=1e-12)
    xn = ((rf - mu_true[:, None]) * torch.rsqrt(var_true[:, None] + 1e-12)).float()
    want = xn @ wf.float().t() + bfold
    scale = float(want.abs().mean())
    err = (o2.float() - want).abs()
    assert float(err.mean()) < 0.02 * scale + 4e-3, (float(err.mean()), scale)
    # (3) consumer: residual normalised on the fly
    o3 = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
    D.linear_bf16_fold(a0, w0, None, "resid", o3, resid=r, r_stats=st, r_gamma=gamma, r_beta=beta)
    want3 = F.layer_norm(r.float(), (N,), gamma, beta, 1e-12)
    err3 = (o3.float() - want3).abs()
    assert float(err3.mean()) < 0.01 * float(want3.abs().mean()) + 2e-3, (float(err3.mean()), float(want3.abs().mean()))


def test_device_side_row_count_and_accumulation():
    ops.require_native()
    g = torch.Generator(device=DEV).manual_seed(2)
    M, K, N = 6000, 768, 768
    a = (torch.randn(M, K, device=DEV, generator=g) * 0.5).bfloat16()
    w = (torch.randn(N, K, device=DEV, generator=g) * 0.03).bfloat16()
    r = torch.randn(M, N, device=DEV, generator=g).bfloat16()
    out = torch.full((M, N), 7.0, device=DEV, dtype=torch.bfloat16)
    st = torch.zeros(M, N // 64, 2, device=DEV)
    cnt = torch.tensor([133], device=DEV, dtype=torch.int32)
    D.linear_bf16_fold(a, w, None, "resid", out, resid=r, stats_out=st, m_dev=cnt, m_scale=30)
    rows = 133 * 30
    want = a[:rows].float() @ w.float().t() + r[:rows].float()
    assert float((out[:rows].float() - want).abs().max()) < 0.06
    assert (out[rows:] == 7.0).all() and (st[rows:] == 0).all()          # nothing past the device-side count is touched
    torch.testing.assert_close(st[:rows], _stats(want), atol=0.08, rtol=3e-3)


def test_folded_trunk_matches_the_unfolded_one():
    """Whole step with cfg.ln_fold on/off: same loss, news vectors and gradients up to bf16 noise."""
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.data import make_synthetic_mind
    from fedrec_b200.engine import TrainStep
    from fedrec_b200.models import UserModel
    news_index = make_synthetic_mind(num_news=4096, num_train=8, num_valid=1, title_len=30, seed=0).news_index
    rng = np.random.default_rng(3)
    cand = torch.from_numpy(rng.integers(1, 4096, size=(16, 5))).to(torch.int32).to(DEV)
    his = torch.from_numpy(rng.integers(1, 4096, size=(16, 50))).to(torch.int32).to(DEV)
    res = {}
    for fold in (False, True):
        torch.manual_seed(0)
        cfg = FedRecConfig(batch_size=16, title_len=30, num_news=4096, cuda_graph=fold, dropout=0.0, ln_fold=fold)
        model = UserModel(None, news_index, DEV, config=cfg)
        with torch.no_grad():                                           # non-trivial LayerNorm affines
            gen = torch.Generator().manual_seed(5)
            for blk in model.text_encoder.DistillBert.transformer.layer:
                for ln in (blk.sa_layer_norm, blk.output_layer_norm):
                    ln.weight.copy_(1.0 + 0.2 * torch.randn(768, generator=gen))
                    ln.bias.copy_(0.1 * torch.randn(768, generator=gen))
        model.on_weights_changed()
        step = TrainStep(model, cfg)
        loss = step.run(cand, his)
        if fold:
            loss = step.run(cand, his)                                  # second call = graph replay
        torch.cuda.synchronize()
        res[fold] = (float(loss), model.flat.grad.clone())
    (l0, g0), (l1, g1) = res[False], res[True]
    assert abs(l0 - l1) < 5e-3, (l0, l1)
    assert float((g0 - g1).norm() / g0.norm()) < 0.08, float((g0 - g1).norm() / g0.norm())
