import pytest

from fedrec_b200.config import FedRecConfig


def test_positional_like_upstream():
    cfg = FedRecConfig.from_argv(["7", "50", "2"], ["total_epochs", "batch_size", "save_every"], mode="grad_avg")
    assert (cfg.total_epochs, cfg.batch_size, cfg.save_every, cfg.mode) == (7, 50, 2, "grad_avg")
    cfg = FedRecConfig.from_argv(["5", "50", "1", "10", "node4"],
                                 ["total_epochs", "batch_size", "save_every", "dp_epsilon", "run_name"])
    assert cfg.dp_epsilon == 10 and cfg.run_name == "node4"


def test_options_and_flags(monkeypatch):
    monkeypatch.setenv("FEDREC_LR", "0.001")
    cfg = FedRecConfig.from_argv(["3", "--clip_norm", "2", "--synthetic", "--no-dedup", "--client-weights=1,2,3"],
                                 ["total_epochs"])
    assert cfg.total_epochs == 3 and cfg.clip_norm == 2.0 and cfg.synthetic and not cfg.dedup
    assert cfg.client_weights == [1.0, 2.0, 3.0] and cfg.lr == 0.001
    with pytest.raises(SystemExit):
        FedRecConfig.from_argv(["--bogus", "1"], [])


def test_kernel_switches(monkeypatch):
    assert FedRecConfig().ln_fold and FedRecConfig().cuda_graph            # the fast paths are the defaults
    assert not FedRecConfig.from_argv(["--ln_fold", "false"], []).ln_fold
    monkeypatch.setenv("FEDREC_LN_FOLD", "0")
    assert not FedRecConfig.from_env().ln_fold
