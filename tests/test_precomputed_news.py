"""BASELINE.json config 1: user encoder (+ the 768->400 projection) trained over precomputed news features.

Upstream anticipates the inference half of this (UserDataset dataset.py:92-109, evaluation_split
evaluation_functions.py:33-46); the training half is the north star's first configuration. Oracles: the ordinary
step on the same weights (the cache is exact at the moment it is built), and the closed-form mean across ranks."""
import os
import socket
import sys

import pytest
import torch
import torch.multiprocessing as mp

from fedrec_b200.config import FedRecConfig
from fedrec_b200.engine import TrainStep
from fedrec_b200.models import UserModel

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _pair(tiny_bert, tiny_data, device="cpu", **kw):
    cfg_a = FedRecConfig(title_len=12, dropout=0.0, batch_size=4, device=str(device), precomputed_news=True, **kw)
    cfg_b = FedRecConfig(title_len=12, dropout=0.0, batch_size=4, device=str(device), **kw)
    torch.manual_seed(0)
    a = UserModel(None, tiny_data.news_index, device, config=cfg_a, bert_config=tiny_bert)
    torch.manual_seed(0)
    b = UserModel(None, tiny_data.news_index, device, config=cfg_b, bert_config=tiny_bert)
    b.load_state_dict(a.state_dict())
    return a, b


def test_precomputed_step_matches_the_full_step_and_never_runs_the_trunk(tiny_bert, tiny_data):
    a, b = _pair(tiny_bert, tiny_data)
    feats = a.build_news_features(batch=64)
    assert feats.shape == (300, 64)

    def boom(*args, **kw):
        raise AssertionError("the trunk must not run in a precomputed_news step")
    a.text_encoder.DistillBert.forward = boom
    g = torch.Generator().manual_seed(1)
    cand = torch.randint(1, 300, (4, 5), generator=g)
    his = torch.randint(0, 300, (4, 50), generator=g)
    sa, sb = TrainStep(a, impl="torch"), TrainStep(b, impl="torch")
    la, lb = sa.run(cand, his), sb.run(cand, his)
    assert float(la) == pytest.approx(float(lb), abs=1e-6)
    for name in a.flat.names:
        ga, gb = a.flat.view(name, grad=True), b.flat.view(name, grad=True)
        if name.startswith("text_encoder.additive_attention"):
            assert float(ga.abs().max()) == 0.0          # frozen into the cache
        else:
            torch.testing.assert_close(ga, gb, atol=1e-6, rtol=1e-4)
    torch.testing.assert_close(sa.scores(cand, his), sb.scores(cand, his), atol=1e-6, rtol=1e-5)
    # training moves the projection and the user encoder
    before = a.flat.param.clone()
    a.adam_step()
    assert not torch.equal(before, a.flat.param)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    torch.set_num_threads(2)
    import torch.distributed as dist
    from fedrec_b200.engine.federated import FederatedRunner
    from fedrec_b200.models.distilbert import DistilBertConfigLite
    bert = DistilBertConfigLite(n_layers=1, dim=32, hidden_dim=64, n_heads=2, hidden_size=32)
    cfg = FedRecConfig(mode="grad_avg", synthetic=True, num_news=128, num_train=32, num_valid=4, title_len=8, batch_size=4,
                       total_epochs=3, steps_per_epoch=2, dropout=0.0, device="cpu", backend="gloo", snapshot_path="",
                       log_dir="", lr=1e-2, precomputed_news=True, heartbeat_s=0.0)
    r = FederatedRunner(cfg, bert_config=bert)
    res = {}
    try:
        r.model.build_news_features()
        r.local_step()
        local = r.model.flat.grad.clone()
        gathered = [torch.zeros_like(local) for _ in range(world)]
        dist.all_gather(gathered, local)
        r.grad_agg.allreduce_mean_()
        res["mean_ok"] = bool(torch.allclose(r.model.flat.grad, sum(gathered) / world, atol=1e-7))
        res["differs"] = bool(not torch.allclose(gathered[0], gathered[1]))
        r.model.adam_step()
        loss = r.run_gradient_averaging()
        p = r.model.flat.param.clone()
        ps = [torch.zeros_like(p) for _ in range(world)]
        dist.all_gather(ps, p)
        res["params_equal"] = bool(torch.equal(ps[0], ps[1]))
        res["loss"] = loss
        res["steps"] = r.model.flat.step
        r.close()
    except Exception as e:  # noqa: BLE001
        import traceback
        res["error"] = f"{e}\n{traceback.format_exc()}"
    q.put((rank, res))


def test_config1_gloo_world2_gradient_averaging_over_precomputed_features():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    out = dict(q.get(timeout=300) for _ in range(2))
    [p.join(30) for p in procs]
    for r in range(2):
        assert "error" not in out[r], out[r].get("error")
        assert out[r]["mean_ok"] and out[r]["differs"] and out[r]["params_equal"]
        assert out[r]["loss"] == out[r]["loss"] and out[r]["steps"] == 1 + 3 * 2


@pytest.mark.gpu
def test_fused_precomputed_step_matches_the_oracle():
    from fedrec_b200 import ops
    from fedrec_b200.data import make_synthetic_mind
    ops.require_native()
    data = make_synthetic_mind(num_news=600, num_train=8, num_valid=2, title_len=30, seed=5)
    dev = "cuda:0"
    cfg = FedRecConfig(title_len=30, dropout=0.0, batch_size=6, precomputed_news=True)
    torch.manual_seed(0)
    a = UserModel(None, data.news_index, dev, config=cfg)
    torch.manual_seed(0)
    b = UserModel(None, data.news_index, dev, config=cfg)
    b.load_state_dict(a.state_dict())
    feats = a.build_news_features(batch=256)                 # through the sm_100a trunk + pooling kernels
    b.set_news_features(feats)
    g = torch.Generator(device=dev).manual_seed(1)
    cand = torch.randint(1, 600, (6, 5), device=dev, generator=g, dtype=torch.int32)
    his = torch.randint(0, 600, (6, 50), device=dev, generator=g, dtype=torch.int32)
    fused, oracle = TrainStep(a, impl="fused"), TrainStep(b, impl="torch")
    for it in range(2):                                      # capture, then replay
        la, lb = fused.run(cand, his), oracle.run(cand, his)
        assert float(la) == pytest.approx(float(lb), abs=5e-3)
        ga, gb = a.flat.grad, b.flat.grad
        rel = float((ga - gb).norm() / gb.norm())
        assert rel < 3e-2, rel
        for name in a.flat.names:
            if name.startswith("text_encoder.additive_attention"):
                assert float(a.flat.view(name, grad=True).abs().max()) == 0.0
    torch.testing.assert_close(fused.scores(cand, his), oracle.scores(cand, his), atol=1e-2, rtol=1e-2)
