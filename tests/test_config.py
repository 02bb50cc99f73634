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


def test_an_epoch_is_a_pass_over_the_shard(tiny_bert, solo_group):
    """total_epochs counts passes over the client's shard like upstream's DataLoader loop (main.py:55-91,135-139):
    ceil(impressions / batch_size) steps per epoch, every step an update; --steps_per_epoch overrides."""
    import torch
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.engine.federated import FederatedRunner
    cfg = FedRecConfig(mode="grad_avg", synthetic=True, num_news=120, num_train=10, num_valid=2, title_len=8, batch_size=4,
                       total_epochs=2, dropout=0.0, device="cpu", backend="gloo", snapshot_path="", log_dir="")
    r = FederatedRunner(cfg, bert_config=tiny_bert)
    assert r.steps_per_epoch() == 3                       # ceil(10 / 4)
    r.run_gradient_averaging()
    assert r.model.flat.step == 2 * 3 and r.round == 6
    cfg2 = FedRecConfig(mode="param_avg", synthetic=True, num_news=120, num_train=10, num_valid=2, title_len=8, batch_size=4,
                        total_epochs=2, steps_per_epoch=1, local_steps=2, dropout=0.0, device="cpu", backend="gloo",
                        snapshot_path="", log_dir="")
    r2 = FederatedRunner(cfg2, bert_config=tiny_bert)
    assert r2.steps_per_epoch() == 1
    r2.run_parameter_averaging()
    assert r2.model.flat.step == 2 * 1 * 2                # epochs x rounds per epoch x local steps
    assert torch.isfinite(r2.model.flat.param).all()


def test_validate_every_adds_ranking_metrics_to_the_epoch_log(tiny_bert, solo_group):
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.engine.federated import FederatedRunner
    cfg = FedRecConfig(mode="grad_avg", synthetic=True, num_news=120, num_train=8, num_valid=6, title_len=8, batch_size=4,
                       total_epochs=3, validate_every=2, dropout=0.0, device="cpu", backend="gloo", snapshot_path="", log_dir="")
    r = FederatedRunner(cfg, bert_config=tiny_bert)
    seen = []
    log = r.logger.log
    r.logger.log = lambda rec, step=None: (seen.append(dict(rec)), log(rec, step=step))[1]
    r.run_gradient_averaging()
    with_val = [x["epoch"] for x in seen if "valid_auc" in x]
    assert with_val == [1] and all(0.0 <= x["valid_auc"] <= 1.0 for x in seen if "valid_auc" in x)
