import math

import pytest

from fedrec_b200.utils.privacy import compute_epsilon, get_noise_multiplier, rdp_subsampled_gaussian


def test_full_batch_rdp_closed_form():
    assert rdp_subsampled_gaussian(1.0, 2.0, 8) == pytest.approx(8 / (2 * 4.0))


def test_epsilon_monotone_in_sigma_and_steps():
    e = [compute_epsilon(s, 0.01, 1000, 1e-5) for s in (0.6, 0.8, 1.0, 1.5, 3.0)]
    assert all(a > b for a, b in zip(e, e[1:]))
    assert compute_epsilon(1.0, 0.01, 2000, 1e-5) > compute_epsilon(1.0, 0.01, 1000, 1e-5)


def test_known_regime():
    # the classic DP-SGD setting (q=256/60000, sigma=1.1, 60 epochs, delta=1e-5) is reported at eps ~ 3.0-3.2
    q = 256 / 60000
    eps = compute_epsilon(1.1, q, int(60 / q), 1e-5)
    assert 2.5 < eps < 3.6


def test_calibration_hits_target():
    sigma = get_noise_multiplier(10.0, 1e-5, sample_rate=0.0125, epochs=50)   # upstream constants client.py:220-224
    eps = compute_epsilon(sigma, 0.0125, math.ceil(50 / 0.0125), 1e-5)
    assert eps <= 10.0 and eps > 9.5
    assert get_noise_multiplier(1.0, 1e-5, 0.0125, epochs=50) > sigma


def test_the_run_log_reports_the_privacy_spent(tiny_bert, solo_group):
    """Clipped + noised gradient averaging logs the (epsilon, delta) its steps have cost so far, per epoch."""
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.engine.federated import FederatedRunner
    from fedrec_b200.utils.privacy import compute_epsilon
    cfg = FedRecConfig(mode="grad_avg", synthetic=True, num_news=120, num_train=40, num_valid=2, title_len=8, batch_size=4,
                       total_epochs=2, steps_per_epoch=3, dropout=0.0, device="cpu", backend="gloo", clip_norm=2.0,
                       noise_multiplier=1.0, snapshot_path="", log_dir="")
    r = FederatedRunner(cfg, bert_config=tiny_bert)
    seen = []
    log = r.logger.log
    r.logger.log = lambda rec, step=None: (seen.append(dict(rec)), log(rec, step=step))[1]
    r.run_gradient_averaging()
    eps = [x["dp_epsilon_spent"] for x in seen if "dp_epsilon_spent" in x]
    assert len(eps) == 2 and eps[0] < eps[1]
    assert abs(eps[1] - compute_epsilon(1.0, 4 / 40, 6, cfg.dp_delta)) < 1e-3
    assert all("impressions_per_s" in x and x["epoch_time_s"] > 0 for x in seen if "epoch" in x)
