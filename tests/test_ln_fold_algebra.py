"""The algebra behind ``fr_gemm_bf16_fold`` in plain fp32 PyTorch (CPU): folding a LayerNorm into the Linear after it."""
import pytest
import torch
import torch.nn.functional as F

from fedrec_b200.ops import reference as ref


@pytest.mark.parametrize("offset,scale", [(0.0, 1.0), (0.7, 1.5), (25.0, 0.5)])
def test_folded_linear_equals_layernorm_then_linear(offset, scale):
    g = torch.Generator().manual_seed(0)
    M, K, N = 64, 768, 96
    x = (torch.randn(M, K, generator=g) * scale + offset).bfloat16()          # what the producing GEMM stored
    x[:, 5] += 40.0                                                            # an outlier feature, as real BERT states have
    w = torch.randn(N, K, generator=g) * 0.03
    b = torch.randn(N, generator=g) * 0.1
    gamma = 1.0 + 0.2 * torch.randn(K, generator=g)
    beta = 0.1 * torch.randn(K, generator=g)
    wf, c, bf = ref.fold_layernorm(w, b, gamma, beta)
    assert wf.dtype == torch.bfloat16 and torch.equal(c, wf.float().sum(1))
    got = ref.linear_after_layernorm_folded(x, wf, c, bf)
    # same rounded operands, normalise first: identical up to fp32 rounding of the re-association
    want_same = F.layer_norm(x.float(), (K,), None, None, 1e-12) @ wf.float().t() + bf
    assert float((got - want_same).abs().max()) < 2e-3
    # the textbook expression with the unrounded weights: only the bf16 rounding of W∘gamma separates the two
    want = F.layer_norm(x.float(), (K,), gamma, beta, 1e-12) @ w.t() + b
    assert float((got - want).abs().max()) < 0.05 and float((got - want).abs().mean()) < 5e-3


def test_constant_rows_cancel_exactly():
    """c is the row sum of the *rounded* W', so a row with no variance maps to b' exactly (0 * rstd is still 0 here)."""
    g = torch.Generator().manual_seed(1)
    w = torch.randn(16, 128, generator=g)
    wf, c, bf = ref.fold_layernorm(w, torch.zeros(16), torch.ones(128), torch.zeros(128))
    x = torch.full((4, 128), 3.0).bfloat16()
    xw = x.float() @ wf.float().t()
    assert torch.allclose(xw - 3.0 * c[None, :], torch.zeros(4, 16), atol=1e-4)


def test_packed_trunk_builds_the_folded_weights_in_place(monkeypatch):
    """PackedTrunk's packing code on the CPU (the bf16 cast kernel is replaced by torch): keys, shapes, the fold values and
    the in-place guarantee that captured graphs rely on."""
    from fedrec_b200.engine import fused
    from fedrec_b200.models.distilbert import DistilBertConfigLite, DistilBertTrunk
    from fedrec_b200.ops import dispatch as D

    def cast(x, out=None):
        return x.bfloat16() if out is None else out.copy_(x)
    monkeypatch.setattr(D, "cast_bf16", cast)
    torch.manual_seed(0)
    bert = DistilBertTrunk(DistilBertConfigLite(n_layers=2, dim=64, hidden_dim=128, n_heads=2))
    with torch.no_grad():
        for blk in bert.transformer.layer:
            blk.sa_layer_norm.weight.uniform_(0.5, 1.5)
            blk.output_layer_norm.bias.uniform_(-0.2, 0.2)
    tr = fused.PackedTrunk(bert)
    L0, L1 = tr.layers
    assert "wqkvf" not in L0 and {"w1f", "c1", "b1f"} <= set(L0) and {"wqkvf", "cqkv", "bqkvf", "w1f"} <= set(L1)
    assert L1["wqkvf"].shape == (3 * 64, 64) and L1["wqkvf"].dtype == torch.bfloat16 and L1["cqkv"].shape == (3 * 64,)
    blk0, blk1 = bert.transformer.layer
    wf, c, bf = ref.fold_layernorm(blk0.ffn.lin1.weight.data, blk0.ffn.lin1.bias.data, blk0.sa_layer_norm.weight.data,
                                   blk0.sa_layer_norm.bias.data)
    assert torch.equal(L0["w1f"], wf) and torch.equal(L0["c1"], c) and torch.equal(L0["b1f"], bf)
    a = blk1.attention
    wq = torch.cat([a.q_lin.weight.data, a.k_lin.weight.data, a.v_lin.weight.data], 0)
    bq = torch.cat([a.q_lin.bias.data, a.k_lin.bias.data, a.v_lin.bias.data])
    wf, c, bf = ref.fold_layernorm(wq, bq, blk0.output_layer_norm.weight.data, blk0.output_layer_norm.bias.data)
    assert torch.equal(L1["wqkvf"], wf) and torch.allclose(L1["bqkvf"], bf)
    ptrs = {k: v.data_ptr() for k, v in L1.items()}
    with torch.no_grad():
        blk0.output_layer_norm.weight.mul_(2.0)
    tr.repack()
    assert {k: v.data_ptr() for k, v in L1.items()} == ptrs                   # refilled in place
    assert not torch.equal(L1["wqkvf"], wf)                                   # and with the new values


def _attention_core(blk, qkv, mask):
    """Soft-max attention of one block from an already projected (q | k | v), then nothing else (no out-projection)."""
    n, T, D3 = qkv.shape
    D = D3 // 3
    H = blk.attention.n_heads
    dh = D // H
    q, k, v = (t.view(n, T, H, dh).transpose(1, 2) for t in qkv.split(D, dim=-1))
    s = (q / dh ** 0.5) @ k.transpose(-1, -2)
    if mask is not None:
        s = s.masked_fill((mask == 0)[:, None, None, :], torch.finfo(s.dtype).min)
    return (torch.softmax(s, -1) @ v).transpose(1, 2).reshape(n, T, D)


def test_folded_trunk_wiring_equals_the_module_forward():
    """The wiring of FusedStep._trunk_folded stated in fp32 PyTorch: un-normalised block outputs travel between the
    GEMMs, FFN1 folds sa_layer_norm of its block, QKV of block l >= 1 folds output_layer_norm of block l-1, residuals are
    normalised on the fly, only the last output_layer_norm is applied as such. Must equal DistilBertTrunk.forward."""
    from fedrec_b200.models.distilbert import DistilBertConfigLite, DistilBertTrunk
    torch.manual_seed(0)
    bert = DistilBertTrunk(DistilBertConfigLite(n_layers=3, dim=64, hidden_dim=128, n_heads=4)).eval()
    with torch.no_grad():
        for blk in bert.transformer.layer:
            for ln in (blk.sa_layer_norm, blk.output_layer_norm):
                ln.weight.uniform_(0.5, 1.5)
                ln.bias.uniform_(-0.2, 0.2)
    ids = torch.randint(1, 1000, (5, 12))
    mask = torch.ones(5, 12, dtype=torch.long)
    mask[:, 9:] = 0
    with torch.no_grad():
        want = bert(ids, mask)[0]
        x = bert.embeddings(ids)                                   # normalised input of block 0
        n, T, D = x.shape
        flat = lambda t: t.reshape(n * T, -1)                      # noqa: E731
        ln = lambda t, m: F.layer_norm(t, (D,), m.weight, m.bias, 1e-12)   # noqa: E731

        def fold_lin(inp, w, b, norm):                             # Linear(LayerNorm(inp)) from the un-normalised inp
            wf, c, bf = ref.fold_layernorm(w, b, norm.weight, norm.bias)
            return ref.linear_after_layernorm_folded(flat(inp), wf.float(), c, bf).view(n, T, -1)

        y2, prev = None, None
        for l, blk in enumerate(bert.transformer.layer):
            a = blk.attention
            wqkv = torch.cat([a.q_lin.weight, a.k_lin.weight, a.v_lin.weight], 0)
            bqkv = torch.cat([a.q_lin.bias, a.k_lin.bias, a.v_lin.bias])
            if l == 0:
                qkv, resid = x @ wqkv.t() + bqkv, x
            else:
                qkv, resid = fold_lin(y2, wqkv, bqkv, prev.output_layer_norm), ln(y2, prev.output_layer_norm)
            y1 = a.out_lin(_attention_core(blk, qkv, mask)) + resid            # un-normalised
            h = F.gelu(fold_lin(y1, blk.ffn.lin1.weight, blk.ffn.lin1.bias, blk.sa_layer_norm))
            y2 = blk.ffn.lin2(h) + ln(y1, blk.sa_layer_norm)                    # un-normalised
            prev = blk
        got = ln(y2, prev.output_layer_norm)
    # fold_layernorm rounds W∘gamma to bf16 (it is the GEMM operand), so the comparison is at bf16 weight precision
    assert float((got - want).abs().max()) < 0.05 and float((got - want).abs().mean()) < 5e-3
