"""User-encoder / loss / weight-gradient kernels vs fp32 PyTorch autograd of the same math."""
import pytest
import torch

from fedrec_b200 import ops
from fedrec_b200.ops import dispatch as D, reference as ref

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("K,M1,N1", [(2500, 400, 400), (2500, 200, 400), (2750, 400, 768), (8250, 384, 768),
                                     (64, 64, 64), (100, 128, 256), (82500, 384, 768)])
def test_gemm_grad_matches_torch(K, M1, N1):
    ops.require_native()
    g = torch.Generator(device=DEV).manual_seed(K + M1)
    A = (torch.randn(K, M1, device=DEV, generator=g) * 0.3).bfloat16()
    B = (torch.randn(K, N1, device=DEV, generator=g) * 0.3).bfloat16()
    out = torch.full((M1, N1), 0.5, device=DEV)
    D.gemm_grad(A, B, out)
    torch.cuda.synchronize()
    want = A.float().t() @ B.float() + 0.5
    tol = 2e-3 * max(1.0, float(want.abs().max()))
    torch.testing.assert_close(out, want, atol=tol, rtol=2e-3)


def test_gemm_grad_dynamic_k_and_strided_operand():
    ops.require_native()
    g = torch.Generator(device=DEV).manual_seed(7)
    K, T = 3000, 30
    A_full = (torch.randn(K, 1200, device=DEV, generator=g) * 0.3).bfloat16()
    B = (torch.randn(K, 400, device=DEV, generator=g) * 0.3).bfloat16()
    cnt = torch.tensor([37], device=DEV, dtype=torch.int32)
    rows = 37 * T
    A_full[rows: (rows + 63) // 64 * 64] = 0            # contract: tail of the last 64-row block is zero
    A = A_full[:, 400:800]                               # column slice: lda = 1200
    out = torch.zeros(400, 400, device=DEV)
    D.gemm_grad(A, B, out, k_dev=cnt, k_scale=T)
    torch.cuda.synchronize()
    want = A[:rows].float().t() @ B[:rows].float()
    torch.testing.assert_close(out, want, atol=2e-2, rtol=2e-3)


def test_colsum_and_transpose_cast():
    ops.require_native()
    x = torch.randn(2500, 1200, device=DEV).bfloat16()
    out = torch.zeros(400, device=DEV)
    D.colsum(x[:, 400:800], out)
    torch.testing.assert_close(out, x[:, 400:800].float().sum(0), atol=1e-2, rtol=1e-3)
    rw = torch.randn(2500, device=DEV)
    cnt = torch.tensor([70], device=DEV, dtype=torch.int32)
    outw = torch.zeros(400, device=DEV)
    D.colsum_weighted(x[:, 400:800], rw, outw, m_dev=cnt, m_scale=30)       # rows limited to 70 * 30 by the device count
    want = (rw[:2100, None] * x[:2100, 400:800].float()).sum(0)
    torch.testing.assert_close(outw, want, atol=2e-2, rtol=2e-3)
    w = torch.randn(200, 400, device=DEV)
    wt = D.transpose_cast(w)
    assert torch.equal(wt, w.t().contiguous().bfloat16())


def _params(seed=0, D_=400, H=20, dk=20, Q=200):
    g = torch.Generator(device=DEV).manual_seed(seed)
    r = lambda *s, sc=0.05: (torch.randn(*s, device=DEV, generator=g) * sc)  # noqa: E731
    return dict(wq=r(H * dk, D_), bq=r(H * dk), wk=r(H * dk, D_), bk=r(H * dk), wv=r(H * dk, D_), bv=r(H * dk),
                w1=r(Q, H * dk), b1=r(Q), w2=r(1, Q, sc=0.2), b2=r(1))


@pytest.mark.parametrize("use_mask", [False, True])
def test_mha_and_pool_score_kernels(use_mask):
    """Kernel-level check on bf16-rounded inputs so the comparison isolates the kernels' own arithmetic."""
    ops.require_native()
    B, L, C, H, dk, Dm, Q = 7, 50, 5, 20, 20, 400, 200
    g = torch.Generator(device=DEV).manual_seed(3)
    qkv = (torch.randn(B * L, 3 * H * dk, device=DEV, generator=g) * 0.8).bfloat16()
    his = torch.randint(0, 3, (B, L), device=DEV, generator=g, dtype=torch.int32)
    his[:, 0] = 1
    ctx = torch.zeros(B * L, H * dk, device=DEV, dtype=torch.bfloat16)
    stats = torch.zeros(B * H * L, 2, device=DEV)
    D.ue_mha_fwd(qkv, his.reshape(-1), ctx, stats, B, L, H, dk, use_mask)
    qf = qkv.float().requires_grad_()
    q, k, v = [t.view(B, L, H, dk).transpose(1, 2) for t in qf.split(H * dk, dim=1)]
    s = q @ k.transpose(-1, -2) / dk ** 0.5
    km = (his != 0).float()[:, None, None, :] if use_mask else None
    want_ctx = (ref.exp_normalise(s, -1, km) @ v).transpose(1, 2).reshape(B * L, H * dk)
    torch.testing.assert_close(ctx.float(), want_ctx, atol=2e-2, rtol=2e-2)
    # backward of the attention alone
    dctx = (torch.randn(B * L, H * dk, device=DEV, generator=g) * 0.1).bfloat16()
    dqkv = torch.zeros_like(qkv)
    D.ue_mha_bwd(qkv, dctx, stats, his.reshape(-1), dqkv, B, L, H, dk, use_mask)
    want_ctx.backward(dctx.float())
    torch.testing.assert_close(dqkv.float(), qf.grad, atol=3e-3, rtol=3e-2)
    # pooling + scoring + CE, forward and backward
    P = _params(1)
    U = 40
    nv = (torch.randn(U, Dm, device=DEV, generator=g) * 0.3).requires_grad_()
    inv_cand = torch.randint(0, U, (B * C,), device=DEV, generator=g, dtype=torch.int32)
    cx = ctx.float().detach().requires_grad_()
    e_pre = (cx @ P["w1"].t() + P["b1"])
    E = torch.tanh(e_pre).bfloat16()
    Ef = E.float().detach().requires_grad_()
    w2 = P["w2"].view(-1).clone().requires_grad_()
    b2 = P["b2"].clone().requires_grad_()
    score = torch.zeros(B, C, device=DEV)
    loss = torch.zeros((), device=DEV)
    user = torch.zeros(B, Dm, device=DEV)
    d_ctx_pool, d_pre = torch.zeros_like(ctx), torch.zeros(B * L, Q, device=DEV, dtype=torch.bfloat16)
    d_w2, d_b2, d_news = torch.zeros(Q, device=DEV), torch.zeros(1, device=DEV), torch.zeros(U, Dm, device=DEV)
    D.ue_pool_score(ctx, E, w2.detach(), b2.detach(), nv.detach(), inv_cand, his.reshape(-1), score, loss, user,
                    d_ctx_pool, d_pre, d_w2, d_b2, d_news, B, L, C, use_mask, True)
    a = (Ef @ w2 + b2).view(B, L, 1)
    pm = (his != 0).float()[:, :, None] if use_mask else None
    al = ref.exp_normalise(a, 1, pm)
    uvec = (cx.view(B, L, -1) * al).sum(1)
    cand = nv[inv_cand.long()].view(B, C, Dm)
    wl, ws = ref.score_ce(cand, uvec)
    wl.backward()
    torch.testing.assert_close(loss, wl.detach(), atol=1e-5, rtol=1e-4)
    torch.testing.assert_close(score, ws.detach(), atol=1e-5, rtol=1e-4)
    torch.testing.assert_close(user, uvec.detach(), atol=1e-5, rtol=1e-4)
    torch.testing.assert_close(d_news, nv.grad, atol=1e-6, rtol=1e-3)
    torch.testing.assert_close(d_w2, w2.grad, atol=1e-6, rtol=2e-3)
    torch.testing.assert_close(d_b2, b2.grad, atol=1e-6, rtol=2e-3)
    torch.testing.assert_close(d_ctx_pool.float(), cx.grad, atol=1e-5, rtol=2e-2)     # pooling path only
    want_dpre = Ef.grad * (1 - Ef.detach() ** 2)
    torch.testing.assert_close(d_pre.float(), want_dpre, atol=1e-6, rtol=2e-2)


def test_dropout_mask_is_consistent_between_gather_and_scatter():
    ops.require_native()
    U, R, Dm = 300, 2500, 400
    g = torch.Generator(device=DEV).manual_seed(5)
    nv = torch.randn(U, Dm, device=DEV, generator=g) + 3.0           # no exact zeros
    inv = torch.randint(0, U, (R,), device=DEV, generator=g, dtype=torch.int32)
    X = torch.zeros(R, Dm, device=DEV, dtype=torch.bfloat16)
    D.ue_gather_dropout(nv, inv, X, 0.2, 1234)
    keep = (X != 0)
    frac = float(keep.float().mean())
    assert frac == pytest.approx(0.8, abs=0.005)
    torch.testing.assert_close(X.float()[keep], (nv[inv.long()] * 1.25)[keep], atol=5e-2, rtol=1e-2)
    dX = torch.ones(R, Dm, device=DEV)
    d_news = torch.zeros(U, Dm, device=DEV)
    D.ue_scatter_dx(dX, inv, d_news, 0.2, 1234)
    want = torch.zeros(U, Dm, device=DEV).index_add_(0, inv.long(), keep.float() * 1.25)
    torch.testing.assert_close(d_news, want, atol=1e-4, rtol=1e-5)
    X2 = torch.zeros_like(X)
    D.ue_gather_dropout(nv, inv, X2, 0.2, 1235)
    assert not torch.equal(X2 != 0, keep)                           # a new step draws a new mask


def test_text_head_backward_matches_autograd():
    ops.require_native()
    n, T, Hq, Hd = 70, 30, 384, 768
    g = torch.Generator(device=DEV).manual_seed(11)
    hidden = torch.randn(n * T, Hd, device=DEV, generator=g).bfloat16()
    w1 = (torch.randn(Hq, Hd, device=DEV, generator=g) * 0.03)
    b1 = torch.randn(Hq, device=DEV, generator=g) * 0.1
    w2 = (torch.randn(Hq, device=DEV, generator=g) * 0.1)
    b2 = torch.randn(1, device=DEV, generator=g) * 0.1
    cnt = torch.tensor([61], device=DEV, dtype=torch.int32)
    e = D.linear_bf16(hidden, w1.bfloat16(), b1, "tanh")
    pooled = torch.zeros(n, Hd, device=DEV, dtype=torch.bfloat16)
    alpha = torch.zeros(n * T, device=DEV)
    D.text_pool(hidden, e, w2, b2, None, None, cnt, n, T, pooled, alpha)
    d_pooled = (torch.randn(n, Hd, device=DEV, generator=g) * 0.1).bfloat16()
    d_pre = torch.full((n * T, Hq), 9.0, device=DEV, dtype=torch.bfloat16)
    d_w2, d_b2, db1 = torch.zeros(Hq, device=DEV), torch.zeros(1, device=DEV), torch.zeros(Hq, device=DEV)
    da = torch.zeros(n * T, device=DEV)
    D.text_pool_bwd(hidden, e, alpha, d_pooled, w2, cnt, n, T, d_pre, da, d_w2, d_b2, db1)
    dW1 = torch.zeros(Hq, Hd, device=DEV)
    D.gemm_grad(d_pre, hidden, dW1, k_dev=cnt, k_scale=T)
    db1_check = torch.zeros(Hq, device=DEV)
    D.colsum(d_pre, db1_check, m_dev=cnt, m_scale=T)
    assert float((db1 - db1_check).norm() / db1_check.norm()) < 1e-3          # d_b1 is the column sum of the stored d_pre
    # split form used by the training step: the two all-row reductions as separate launches (side stream there)
    d_pre2 = torch.full_like(d_pre, 9.0)
    d_w2s, d_b2s, db1s, da2 = torch.zeros(Hq, device=DEV), torch.zeros(1, device=DEV), torch.zeros(Hq, device=DEV), torch.zeros(n * T, device=DEV)
    D.text_pool_bwd(hidden, e, alpha, d_pooled, w2, cnt, n, T, d_pre2, da2, None, d_b2s, None)
    D.colsum_weighted(e, da2, d_w2s, m_dev=cnt, m_scale=T)
    D.colsum(d_pre2, db1s, m_dev=cnt, m_scale=T)
    assert torch.equal(d_pre2, d_pre) and torch.equal(da2, da)
    torch.testing.assert_close(d_w2s, d_w2, atol=1e-4, rtol=1e-3)
    torch.testing.assert_close(db1s, db1, atol=1e-4, rtol=1e-3)
    torch.cuda.synchronize()
    m = 61
    h = hidden[: m * T].float().view(m, T, Hd)
    w1r = w1.bfloat16().float().requires_grad_()
    b1r, w2r, b2r = b1.clone().requires_grad_(), w2.clone().requires_grad_(), b2.clone().requires_grad_()
    pooled_ref = ref.additive_pool(h, w1r, b1r, w2r.view(1, -1), b2r)
    torch.testing.assert_close(pooled[:m].float(), pooled_ref.detach(), atol=2e-2, rtol=2e-2)
    pooled_ref.backward(d_pooled[:m].float())
    assert (d_pre[m * T: (m * T + 63) // 64 * 64] == 0).all() and (d_pre[(m * T + 63) // 64 * 64:] == 9.0).all()
    rel = lambda a, b_: float((a - b_).norm() / b_.norm())  # noqa: E731
    assert rel(d_w2, w2r.grad) < 3e-2
    # d_b2 = sum_t d_a_t = dot * (1 - sum alpha) is ~0 by construction: compare on the scale of the other gradients
    assert abs(float(d_b2) - float(b2r.grad)) < 1e-3 * float(w2r.grad.abs().max())
    assert rel(dW1, w1r.grad) < 3e-2, rel(dW1, w1r.grad)
    assert rel(db1, b1r.grad) < 3e-2


@pytest.mark.parametrize("B,L,H,dk,use_mask", [(7, 50, 20, 20, False), (7, 50, 20, 20, True), (50, 50, 20, 20, False),
                                               (3, 64, 4, 32, True), (5, 33, 6, 16, False), (1, 1, 2, 20, False)])
def test_tensor_core_user_attention_matches_the_reference_and_the_cuda_core_kernel(B, L, H, dk, use_mask):
    """fr_ue_attention_tc (tcgen05, two users per 128-row tile, head dim padded to 32) against attention.py's
    exp(s)/(sum + 1e-8) attention in fp32 and against the CUDA-core kernel, including the (max, denominator) pairs the
    backward kernel consumes; the backward run on the tensor-core kernel's statistics must match autograd."""
    ops.require_native()
    g = torch.Generator(device=DEV).manual_seed(11)
    qkv = (torch.randn(B * L, 3 * H * dk, device=DEV, generator=g) * 0.8).bfloat16()
    his = torch.randint(0, 3, (B, L), device=DEV, generator=g, dtype=torch.int32)
    his[:, 0] = 1
    ctx_tc = torch.full((B * L, H * dk), 7.0, device=DEV, dtype=torch.bfloat16)
    ctx_cc = torch.zeros_like(ctx_tc)
    st_tc = torch.zeros(B * H * L, 2, device=DEV)
    st_cc = torch.zeros_like(st_tc)
    D.ue_mha_fwd(qkv, his.reshape(-1), ctx_tc, st_tc, B, L, H, dk, use_mask, impl="tc")
    D.ue_mha_fwd(qkv, his.reshape(-1), ctx_cc, st_cc, B, L, H, dk, use_mask, impl="cuda")
    qf = qkv.float().requires_grad_()
    q, k, v = [t.view(B, L, H, dk).transpose(1, 2) for t in qf.split(H * dk, dim=1)]
    s = q @ k.transpose(-1, -2) / dk ** 0.5
    km = (his != 0).float()[:, None, None, :] if use_mask else None
    want = (ref.exp_normalise(s, -1, km) @ v).transpose(1, 2).reshape(B * L, H * dk)
    torch.testing.assert_close(ctx_tc.float(), want, atol=2e-2, rtol=2e-2)
    torch.testing.assert_close(ctx_tc.float(), ctx_cc.float(), atol=2e-2, rtol=2e-2)
    torch.testing.assert_close(st_tc[:, 0], st_cc[:, 0], atol=2e-3, rtol=1e-4)       # row maxima (natural-log units)
    torch.testing.assert_close(st_tc[:, 1], st_cc[:, 1], atol=1e-4, rtol=2e-3)       # denominators
    dctx = (torch.randn(B * L, H * dk, device=DEV, generator=g) * 0.1).bfloat16()
    want.backward(dctx.float())
    for impl in ("tc", "cuda"):
        dqkv = torch.full_like(qkv, 7.0)
        D.ue_mha_bwd(qkv, dctx, st_tc, his.reshape(-1), dqkv, B, L, H, dk, use_mask, impl=impl)
        # bf16 probabilities / dS on the tensor cores: compare in the norm per q|k|v block plus a loose element bound
        for o, nm in enumerate("qkv"):
            a, b = dqkv.float()[:, o * H * dk:(o + 1) * H * dk], qf.grad[:, o * H * dk:(o + 1) * H * dk]
            if float(b.norm()) < 1e-6:          # a history of one click: the soft-max is constant, dQ = dK = 0
                assert float(a.abs().max()) < 1e-4
                continue
            rel = float((a - b).norm() / b.norm())
            assert rel < (2e-2 if impl == "tc" else 1e-2), (impl, nm, rel, (B, L, H, dk, use_mask))
        torch.testing.assert_close(dqkv.float(), qf.grad, atol=6e-3 if impl == "tc" else 3e-3, rtol=5e-2)
