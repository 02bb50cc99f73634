"""CPU semantics of the aggregation wrappers (single process): clip bound, noise scale, weighted FedAvg edge cases."""
import pytest
import torch

from fedrec_b200 import ops
from fedrec_b200.parallel.aggregate import GradAggregator, ParamAggregator


class _Flat:
    def __init__(self, n):
        self.grad = torch.zeros(n)


def test_clip_bounds_the_flat_gradient_norm():
    flat = _Flat(4096)
    torch.manual_seed(0)
    flat.grad.copy_(torch.randn(4096))
    flat.grad.mul_(10.0 / flat.grad.norm())
    agg = GradAggregator(flat, 1, 0, clip_norm=2.0, noise_multiplier=0.0)
    agg.allreduce_mean_()
    assert abs(float(flat.grad.norm()) - 2.0) < 1e-4
    small = torch.full((16,), 0.01)
    before = small.clone()
    agg.allreduce_mean_(small)                       # already inside the ball: untouched
    assert torch.equal(small, before)


def test_noise_is_sigma_times_clip_and_differs_per_round_and_rank():
    n = 200_000
    flat = _Flat(n)
    agg = GradAggregator(flat, 1, 0, clip_norm=2.0, noise_multiplier=0.05, seed=3)
    assert agg.noise_std == pytest.approx(0.1)       # DP-SGD convention: sigma is relative to the clip bound
    agg.allreduce_mean_()
    a = flat.grad.clone()
    assert float(a.std()) == pytest.approx(0.1, rel=0.02) and abs(float(a.mean())) < 2e-3
    flat.grad.zero_()
    agg.allreduce_mean_()
    assert not torch.equal(a, flat.grad)             # a new round draws new noise
    x0, x1 = torch.zeros(1000), torch.zeros(1000)
    ops.clip_noise_(x0, 0.0, 0.1, seed=11, stream_id=0)
    ops.clip_noise_(x1, 0.0, 0.1, seed=11, stream_id=1)
    assert not torch.equal(x0, x1)                   # clients draw independent streams from the same seed
    y0 = torch.zeros(1000)
    ops.clip_noise_(y0, 0.0, 0.1, seed=11, stream_id=0)
    assert torch.equal(x0, y0)                       # and the stream is reproducible
    assert GradAggregator(flat, 1, 0, clip_norm=0.0, noise_multiplier=0.05).noise_std == pytest.approx(0.05)


def test_param_average_single_rank_and_bad_weights():
    buf = torch.arange(8, dtype=torch.float32)
    agg = ParamAggregator(1, 0)
    assert torch.equal(agg.average_(buf.clone()), buf)
    with pytest.raises(ValueError):
        ParamAggregator(2, 0).average_(buf.clone(), [0.0, 0.0])
