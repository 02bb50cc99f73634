"""The fused sm_100a step vs the fp32 autograd oracle, same weights, same batch."""
import pytest
import torch

from fedrec_b200 import ops
from fedrec_b200.config import FedRecConfig
from fedrec_b200.data import make_synthetic_mind
from fedrec_b200.engine import TrainStep
from fedrec_b200.models import UserModel

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _models():
    ops.require_native()
    data = make_synthetic_mind(num_news=600, num_train=8, num_valid=2, title_len=30, seed=5)
    cfg = FedRecConfig(title_len=30, dropout=0.0, batch_size=6)
    torch.manual_seed(0)
    a = UserModel(None, data.news_index, DEV, config=cfg)
    torch.manual_seed(0)
    b = UserModel(None, data.news_index, DEV, config=cfg)
    b.load_state_dict(a.state_dict())
    return a, b


def test_fused_step_matches_oracle():
    a, b = _models()
    g = torch.Generator(device=DEV).manual_seed(1)
    cand = torch.randint(1, 600, (6, 5), device=DEV, generator=g, dtype=torch.int32)
    his = torch.randint(0, 600, (6, 50), device=DEV, generator=g, dtype=torch.int32)
    fused, oracle = TrainStep(a, impl="fused"), TrainStep(b, impl="torch")
    la, lb = fused.run(cand, his), oracle.run(cand, his)
    assert float(la) == pytest.approx(float(lb), abs=5e-3)
    ga, gb = a.flat.grad, b.flat.grad
    assert float(gb.norm()) > 0
    rel = float((ga - gb).norm() / gb.norm())
    assert rel < 5e-2, rel
    for name in a.flat.names:
        x, y = a.flat.view(name, grad=True), b.flat.view(name, grad=True)
        if float(y.norm()) > 1e-8:
            assert float((x - y).norm() / y.norm()) < 0.1, name
    sa, sb = fused.scores(cand, his), oracle.scores(cand, his)
    torch.testing.assert_close(sa, sb, atol=1e-2, rtol=1e-2)


def test_few_steps_track_oracle_loss():
    a, b = _models()
    a.cfg.lr = b.cfg.lr = 1e-3
    fused, oracle = TrainStep(a, impl="fused"), TrainStep(b, impl="torch")
    g = torch.Generator(device=DEV).manual_seed(2)
    for _ in range(5):
        cand = torch.randint(1, 600, (6, 5), device=DEV, generator=g, dtype=torch.int32)
        his = torch.randint(0, 600, (6, 50), device=DEV, generator=g, dtype=torch.int32)
        la, lb = fused.run(cand, his), oracle.run(cand, his)
        a.adam_step(); b.adam_step()
        assert float(la) == pytest.approx(float(lb), abs=1e-2)
