import math

import pytest
import torch

from fedrec_b200 import ops
from fedrec_b200.ops import dispatch as D, reference as ref

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    ops.require_native()
    return torch.device("cuda", 0)


@pytest.mark.parametrize("B,C,D_", [(50, 5, 400), (7, 5, 400), (3, 9, 64)])
def test_score_ce_forward_backward(dev, B, C, D_):
    torch.manual_seed(0)
    cand = (torch.randn(B, C, D_, device=dev) * 0.2).requires_grad_()
    user = (torch.randn(B, D_, device=dev) * 0.2).requires_grad_()
    loss, score = ops.score_ce(cand, user)
    (loss * 1.7).backward()
    c2, u2 = cand.detach().clone().requires_grad_(), user.detach().clone().requires_grad_()
    l2, s2 = ref.score_ce(c2, u2)
    (l2 * 1.7).backward()
    torch.testing.assert_close(loss, l2, atol=1e-5, rtol=1e-5)
    torch.testing.assert_close(score, s2, atol=1e-5, rtol=1e-5)
    torch.testing.assert_close(cand.grad, c2.grad, atol=1e-6, rtol=1e-4)
    torch.testing.assert_close(user.grad, u2.grad, atol=1e-6, rtol=1e-4)


def test_segment_sum(dev):
    torch.manual_seed(1)
    vals = torch.randn(2750, 400, device=dev)
    seg = torch.randint(0, 900, (2750,), device=dev)
    torch.testing.assert_close(ops.segment_sum(vals, seg, 900), ref.segment_sum(vals, seg, 900), atol=1e-4, rtol=1e-5)


def test_dedup_matches_torch_unique(dev):
    torch.manual_seed(2)
    for K, N in [(2750, 65536), (100, 300), (5000, 1_000_003)]:
        ids = torch.randint(0, min(N, 4000), (K,), device=dev, dtype=torch.int32)
        uniq, count, inv = D.dedup(ids, N)
        u, i = torch.unique(ids.long(), return_inverse=True)
        n = int(count.item())
        assert n == u.numel()
        assert torch.equal(uniq[:n].long(), u) and torch.equal(inv.long(), i)
        assert (uniq[n:] == 0).all()
        assert torch.equal(uniq.long()[inv.long()], ids.long())


def test_adam_matches_torch(dev):
    torch.manual_seed(3)
    n = 1_164_896
    p = torch.randn(n, device=dev)
    ref_p = torch.nn.Parameter(p.clone())
    opt = torch.optim.Adam([ref_p], lr=5e-5)

    class Flat:
        pass
    f = Flat()
    f.param, f.grad = p, torch.zeros_like(p)
    f.exp_avg, f.exp_avg_sq, f.step, f.numel = torch.zeros_like(p), torch.zeros_like(p), 0, n
    for _ in range(3):
        g = torch.randn(n, device=dev) * 0.01
        f.grad.copy_(g)
        ref_p.grad = g.clone()
        opt.step()
        ops.adam_flat(f, 5e-5, 0.9, 0.999, 1e-8)
    torch.testing.assert_close(f.param, ref_p.data, atol=1e-6, rtol=1e-5)


def test_clip_and_noise(dev):
    torch.manual_seed(4)
    n = 1_164_896
    g = torch.randn(n, device=dev) * 0.01
    want = float((g.double() ** 2).sum())
    assert float(ops.grad_sq_norm(g)) == pytest.approx(want, rel=1e-4)
    x = g.clone()
    ops.clip_noise_(x, clip=2.0, noise_std=0.0)
    assert float(x.norm()) == pytest.approx(min(2.0, math.sqrt(want)), rel=1e-4)
    torch.testing.assert_close(x, g * min(1.0, 2.0 / math.sqrt(want)), atol=1e-7, rtol=1e-4)
    small = g * 0.01
    y = small.clone()
    ops.clip_noise_(y, clip=2.0, noise_std=0.0)
    torch.testing.assert_close(y, small)                       # inside the ball: untouched
    z = torch.zeros(n, device=dev)
    ops.clip_noise_(z, clip=0.0, noise_std=0.1, seed=99, stream_id=3)
    assert abs(float(z.mean())) < 5e-4 and float(z.std()) == pytest.approx(0.1, rel=0.01)
    kurt = float(((z / z.std()) ** 4).mean())
    assert 2.9 < kurt < 3.1                                       # Gaussian, not uniform
    z2 = torch.zeros(n, device=dev)
    ops.clip_noise_(z2, clip=0.0, noise_std=0.1, seed=99, stream_id=3)
    assert torch.equal(z, z2)                                     # counter-based: reproducible
    z3 = torch.zeros(n, device=dev)
    ops.clip_noise_(z3, clip=0.0, noise_std=0.1, seed=99, stream_id=4)
    assert abs(float((z * z3).mean())) < 1e-4                     # another client's stream is independent


def test_c_host_program_runs():
    """examples/c_abi_demo.c: a plain C host drives the tcgen05 GEMM and the no-allocation de-dup through the C ABI."""
    import subprocess
    from test_build import build_c_demo
    r = subprocess.run([build_c_demo()], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "c_abi_demo: OK" in r.stdout
