"""Noise-multiplier calibration for the Gaussian mechanism, without Opacus.

The reference only uses Opacus as a sigma calculator: it calls
``PrivacyEngine.make_private_with_epsilon`` and keeps ``optimizer.noise_multiplier``
(client.py:269-281, constants :220-224).  Opacus is not installed here and is not
a dependency of this framework, so this module carries a small Rényi-DP accountant
for the Poisson-subsampled Gaussian mechanism (Mironov, Talwar, Zhang 2019) and a
bisection on sigma.  Integer orders only, which gives a valid (slightly
conservative) epsilon.
"""
from __future__ import annotations

import math
from typing import Iterable, Sequence, Tuple

DEFAULT_ORDERS: Tuple[int, ...] = tuple(range(2, 64)) + (72, 80, 96, 128, 192, 256, 384, 512)


def _log_add(a: float, b: float) -> float:
    if a == -math.inf:
        return b
    if b == -math.inf:
        return a
    hi, lo = (a, b) if a > b else (b, a)
    return hi + math.log1p(math.exp(lo - hi))


def _log_comb(n: int, k: int) -> float:
    return math.lgamma(n + 1) - math.lgamma(k + 1) - math.lgamma(n - k + 1)


def rdp_subsampled_gaussian(q: float, sigma: float, alpha: int) -> float:
    """RDP of order ``alpha`` (integer ≥ 2) for one step with sampling rate ``q``."""
    if q == 0:
        return 0.0
    if sigma == 0:
        return math.inf
    if q == 1.0:
        return alpha / (2 * sigma ** 2)
    log_a = -math.inf
    for i in range(alpha + 1):
        term = _log_comb(alpha, i) + i * math.log(q) + (alpha - i) * math.log1p(-q) \
            + (i * i - i) / (2 * sigma ** 2)
        log_a = _log_add(log_a, term)
    return log_a / (alpha - 1)


def epsilon_from_rdp(rdp: Sequence[float], orders: Sequence[int], delta: float) -> Tuple[float, int]:
    best, best_order = math.inf, orders[0]
    for r, a in zip(rdp, orders):
        if a <= 1 or math.isinf(r):
            continue
        eps = r - (math.log(delta) + math.log(a)) / (a - 1) + math.log((a - 1) / a)
        if eps < best:
            best, best_order = eps, a
    return max(best, 0.0), best_order


def compute_epsilon(sigma: float, sample_rate: float, steps: int, delta: float,
                    orders: Iterable[int] = DEFAULT_ORDERS) -> float:
    orders = list(orders)
    rdp = [steps * rdp_subsampled_gaussian(sample_rate, sigma, a) for a in orders]
    return epsilon_from_rdp(rdp, orders, delta)[0]


def get_noise_multiplier(target_epsilon: float, target_delta: float, sample_rate: float,
                         epochs: int = None, steps: int = None, epsilon_tolerance: float = 0.01,
                         orders: Iterable[int] = DEFAULT_ORDERS) -> float:
    """Smallest sigma whose accumulated epsilon after ``steps`` is ≤ ``target_epsilon``."""
    if steps is None:
        if epochs is None:
            raise ValueError("give epochs or steps")
        steps = int(math.ceil(epochs / max(sample_rate, 1e-12)))
    orders = list(orders)
    lo, hi = 0.0, 10.0
    while compute_epsilon(hi, sample_rate, steps, target_delta, orders) > target_epsilon:
        hi *= 2
        if hi > 1e6:
            raise ValueError("target epsilon is unreachable")
    for _ in range(100):
        mid = 0.5 * (lo + hi)
        eps = compute_epsilon(mid, sample_rate, steps, target_delta, orders)
        if eps > target_epsilon:
            lo = mid
        else:
            hi = mid
            if target_epsilon - eps < epsilon_tolerance and hi - lo < 1e-3 * hi:
                break
    return hi
