"""The aggregation chain with all logical ranks emulated on one GPU (blockIdx.y slices of every launch of the chain).

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


@pytest.mark.parametrize("W", [2, 4])
def test_gated_chain_uses_the_device_step_counter_and_runs_once(heaps8, W):
    """The chain a CUDA graph replays: epoch and Adam step count come from the pad; it runs only when gradients are pending."""
    hs = heaps8[W]
    bufs = [getattr(h, "_g", None) if getattr(h, "_g", None) is not None else h.alloc_f32(N) for h in hs]
    dev = bufs[0].device
    torch.manual_seed(40 + W)
    params = [torch.randn(N, device=dev) for _ in range(W)]
    ms = [torch.zeros(N, device=dev) for _ in range(W)]
    vs = [torch.zeros(N, device=dev) for _ in range(W)]
    refs = [torch.nn.Parameter(p.clone()) for p in params]
    opts = [torch.optim.Adam([r], lr=5e-5) for r in refs]
    hs[0].set_step(0)
    kw = dict(clip=0.0, noise_std=0.0, seed=1, scale=1.0 / W, params=params, exp_avg=ms, exp_avg_sq=vs, step=0, lr=5e-5,
              b1=0.9, b2=0.999, eps=1e-8, ranks_here=W, gated=True)
    for step in (1, 2, 3):
        for h, b in zip(hs, bufs):
            h._g = b
            b.copy_(torch.randn(N, device=dev) * 0.01 * (1 + h.rank))
        want = torch.stack([b.clone() for b in bufs]).sum(0) / W
        hs[0].mark_pending(ranks_here=W)
        hs[0].grad_allreduce_adam_(bufs[0], **kw)
        before = [p.clone() for p in params]
        hs[0].grad_allreduce_adam_(bufs[0], **kw)              # nothing pending: must not touch anything
        torch.cuda.synchronize()
        hs[0].check()
        for p_, b_ in zip(params, before):
            assert torch.equal(p_, b_)
        for r, o in zip(refs, opts):
            r.grad = want.clone()
            o.step()
        for b in bufs:
            torch.testing.assert_close(b, want, atol=1e-7, rtol=1e-5)
        assert hs[0].pad_words(73, 1)[0] == step                # device-side step counter
        assert hs[0].pad_words(72, 1)[0] == 0                   # pending cleared
    for p_, r in zip(params, refs):
        torch.testing.assert_close(p_, r.data, atol=1e-6, rtol=1e-5)


def test_single_rank_chain_clips_noises_and_steps():
    """W = 1 (one client on one GPU): the same chain does clip + noise + Adam locally."""
    from fedrec_b200.parallel.symm import SymmetricHeap as SH
    dev = torch.device("cuda", torch.cuda.current_device())
    h = SH.single(dev)
    torch.manual_seed(3)
    g = torch.randn(N, device=dev) * 0.01
    p = torch.randn(N, device=dev)
    m, v = torch.zeros(N, device=dev), torch.zeros(N, device=dev)
    ref = torch.nn.Parameter(p.clone())
    opt = torch.optim.Adam([ref], lr=5e-5)
    clip = 2.0
    sc = min(1.0, clip / float(g.norm()))
    assert sc < 1.0
    want_g = g * sc
    h.grad_allreduce_adam_(g, clip=clip, noise_std=0.0, seed=9, scale=1.0, params=p, exp_avg=m, exp_avg_sq=v, step=1,
                           lr=5e-5, b1=0.9, b2=0.999, eps=1e-8)
    torch.cuda.synchronize()
    torch.testing.assert_close(g, want_g, atol=1e-7, rtol=1e-4)
    ref.grad = want_g.clone()
    opt.step()
    torch.testing.assert_close(p, ref.data, atol=1e-6, rtol=1e-5)
    g2 = torch.zeros(N, device=dev)
    h.grad_allreduce_(g2, clip=0.0, noise_std=0.1, seed=9, scale=1.0)
    torch.cuda.synchronize()
    assert float(g2.std()) == pytest.approx(0.1, rel=0.02) and abs(float(g2.mean())) < 1e-3


def test_a_silent_peer_is_reported_not_trapped():
    """Rank 0 aggregates alone on a 2-rank pad set: rank 1 never arrives. The wait kernel must give up after the
    deadline, name the peer in a host-readable record, leave the CUDA context usable, and ``reset`` must restore the pads."""
    from fedrec_b200.parallel.symm import AggregationError, SymmetricHeap as SH
    hs = SH.local(2, heap_bytes=1 << 20, timeout_s=0.3)
    bufs = [h.alloc_f32(4096) for h in hs]
    for h, b in zip(hs, bufs):
        b.fill_(float(h.rank + 1))
    hs[0].grad_allreduce_(bufs[0], clip=0.0, noise_std=0.0, seed=1, scale=0.5)        # ranks_here=1: only rank 0 runs
    with pytest.raises(AggregationError) as ei:
        hs[0].check(sync=True)
    assert ei.value.rank == 0 and ei.value.peer == 1 and ei.value.status == 1 and ei.value.waited_ms >= 300
    assert "rank 1" in str(ei.value)
    assert torch.equal(bufs[0], torch.full_like(bufs[0], 1.0))                         # the chain behind the wait did nothing
    hs[0].reset(collective=False)
    assert not hs[0].failed()
    hs[0].grad_allreduce_(bufs[0], clip=0.0, noise_std=0.0, seed=1, scale=0.5, ranks_here=2)
    hs[0].check(sync=True)
    for b in bufs:
        assert torch.equal(b, torch.full_like(b, 1.5))
