"""Aggregation kernels, all logical ranks emulated inside one cooperative launch on one GPU.

Oracle: the closed-form mean of the per-rank tensors computed with PyTorch (what an NCCL all-reduce
followed by a divide returns)."""
import math

import pytest
import torch

from fedrec_b200 import ops
from fedrec_b200.parallel.symm import SymmetricHeap

pytestmark = pytest.mark.gpu
N = 1_164_896        # the real flat gradient size (padded)


@pytest.fixture(scope="module")
def heaps8():
    ops.require_native()
    return {w: SymmetricHeap.local(w, heap_bytes=12 << 20) for w in (2, 4, 8)}


@pytest.mark.parametrize("W", [2, 4, 8])
def test_grad_mean_plain(heaps8, W):
    hs = heaps8[W]
    torch.manual_seed(W)
    bufs = [h.alloc_f32(N) for h in hs] if not hasattr(hs[0], "_g") else [h._g for h in hs]
    for h, b in zip(hs, bufs):
        h._g = b
        b.copy_(torch.randn(N, device=b.device) * 0.01 * (1 + h.rank))
    want = torch.stack([b.clone() for b in bufs]).sum(0) / W
    hs[0].grad_allreduce_(bufs[0], clip=0.0, noise_std=0.0, seed=1, scale=1.0 / W, ranks_here=W)
    torch.cuda.synchronize()
    for b in bufs:
        torch.testing.assert_close(b, want, atol=1e-7, rtol=1e-5)


@pytest.mark.parametrize("W", [2, 8])
def test_grad_clip_exact_and_noise_statistics(heaps8, W):
    hs = heaps8[W]
    bufs = [getattr(h, "_g", None) if getattr(h, "_g", None) is not None else h.alloc_f32(N) for h in hs]
    torch.manual_seed(10 + W)
    for h, b in zip(hs, bufs):
        h._g = b
        b.copy_(torch.randn(N, device=b.device) * (0.0005 if h.rank % 2 else 0.01))
    clip = 2.0
    src = [b.clone() for b in bufs]
    scales = [min(1.0, clip / float(s.norm())) for s in src]
    assert any(s < 1 for s in scales) and any(s == 1 for s in scales)
    want = sum(s * k for s, k in zip(src, scales)) / W
    hs[0].grad_allreduce_(bufs[0], clip=clip, noise_std=0.0, seed=5, scale=1.0 / W, ranks_here=W)
    torch.cuda.synchronize()
    for b in bufs:
        torch.testing.assert_close(b, want, atol=1e-7, rtol=1e-4)
    # second call on the same pads (epoch advances, sum-of-squares slot was reset) with noise on
    for b, s in zip(bufs, src):
        b.copy_(s)
    std = 0.05 * clip
    hs[0].grad_allreduce_(bufs[0], clip=clip, noise_std=std, seed=6, scale=1.0 / W, ranks_here=W)
    torch.cuda.synchronize()
    noise = bufs[0] - want
    assert torch.equal(bufs[0], bufs[-1])                                  # every rank holds the same result
    assert abs(float(noise.mean())) < 5 * std / math.sqrt(W * N)
    assert float(noise.std()) == pytest.approx(std / math.sqrt(W), rel=0.02)   # mean of W independent draws


@pytest.mark.parametrize("W", [2, 4, 8])
def test_param_average_weighted_and_broadcast(heaps8, W):
    hs = heaps8[W]
    bufs = [getattr(h, "_g", None) if getattr(h, "_g", None) is not None else h.alloc_f32(N) for h in hs]
    torch.manual_seed(20 + W)
    for h, b in zip(hs, bufs):
        h._g = b
        b.copy_(torch.randn(N, device=b.device))
    src = [b.clone() for b in bufs]
    w = [float(i + 1) for i in range(W)]
    w[1] = 0.0                                                              # an aggregation-only server rank
    tot = sum(w)
    want = sum(s * (x / tot) for s, x in zip(src, w))
    hs[0].param_average_(bufs[0], [x / tot for x in w], ranks_here=W)
    torch.cuda.synchronize()
    for b in bufs:
        torch.testing.assert_close(b, want, atol=1e-6, rtol=1e-5)
    onehot = [0.0] * W
    onehot[W - 1] = 1.0                                                     # broadcast from the last rank
    for b, s in zip(bufs, src):
        b.copy_(s)
    hs[0].param_average_(bufs[0], onehot, ranks_here=W)
    torch.cuda.synchronize()
    for b in bufs:
        assert torch.equal(b, src[W - 1])


@pytest.mark.parametrize("W", [2, 8])
def test_grad_mean_with_fused_adam_tail(heaps8, W):
    """Aggregation + Adam in one kernel == aggregation, then torch.optim.Adam on the averaged gradient."""
    hs = heaps8[W]
    bufs = [getattr(h, "_g", None) if getattr(h, "_g", None) is not None else h.alloc_f32(N) for h in hs]
    torch.manual_seed(30 + W)
    dev = bufs[0].device
    params = [torch.randn(N, device=dev) for _ in range(W)]
    ms = [torch.zeros(N, device=dev) for _ in range(W)]
    vs = [torch.zeros(N, device=dev) for _ in range(W)]
    refs = [torch.nn.Parameter(p.clone()) for p in params]
    opts = [torch.optim.Adam([r], lr=5e-5) for r in refs]
    for step in (1, 2, 3):
        for h, b in zip(hs, bufs):
            h._g = b
            b.copy_(torch.randn(N, device=dev) * 0.01 * (1 + h.rank))
        want = torch.stack([b.clone() for b in bufs]).sum(0) / W
        hs[0].grad_allreduce_adam_(bufs[0], clip=0.0, noise_std=0.0, seed=1, scale=1.0 / W, params=params, exp_avg=ms,
                                   exp_avg_sq=vs, step=step, lr=5e-5, b1=0.9, b2=0.999, eps=1e-8, ranks_here=W)
        torch.cuda.synchronize()
        for r, o in zip(refs, opts):
            r.grad = want.clone()
            o.step()
        for b in bufs:
            torch.testing.assert_close(b, want, atol=1e-7, rtol=1e-5)
    for p_, r in zip(params, refs):
        torch.testing.assert_close(p_, r.data, atol=1e-6, rtol=1e-5)
