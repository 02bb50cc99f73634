"""The one-pass autograd step equals the upstream two-phase flow (leaf grads → segment-sum → head backward)."""
import copy

import numpy as np
import torch

from fedrec_b200.config import FedRecConfig
from fedrec_b200.data import BatchSampler, ImpressionTable
from fedrec_b200.engine import TrainStep, process_news_grad, process_user_grad, train_on_step, Trainer
from fedrec_b200.models import UserModel


def _model(tiny_bert, tiny_data, **kw):
    cfg = FedRecConfig(dropout=0.0, **kw)
    torch.manual_seed(0)
    return UserModel(None, tiny_data.news_index, "cpu", config=cfg, bert_config=tiny_bert)


def _batch(tiny_data, B=6, seed=0):
    tab = ImpressionTable(tiny_data.train_sam, tiny_data.nid2index)
    cand, his = BatchSampler(tab, B, seed=seed, pin=False).next()
    return cand.clone(), his.clone()


def test_one_pass_equals_two_phase(tiny_bert, tiny_data):
    cand, his = _batch(tiny_data)
    a = _model(tiny_bert, tiny_data)
    b = _model(tiny_bert, tiny_data)
    b.load_state_dict(a.state_dict())
    # (a) engine step
    loss_a = TrainStep(a, impl="torch").run(cand, his)
    # (b) upstream-style flow through the compatibility API
    b.train()
    loss_b, score, cand_vecs, his_vecs = b(cand, his, torch.zeros(cand.shape[0], dtype=torch.long))
    b.flat.zero_grad()
    loss_b.backward()
    news_grad = process_news_grad([cand.numpy(), None, cand_vecs.grad], [his.numpy(), None, his_vecs.grad])
    b.collect(news_grad, process_user_grad(b.user_encoder.named_parameters()))
    b.update_news_grad()
    torch.testing.assert_close(loss_a, loss_b.detach(), atol=1e-6, rtol=1e-5)
    torch.testing.assert_close(a.flat.grad, b.flat.grad, atol=2e-6, rtol=1e-4)
    assert a.flat.grad.abs().sum() > 0


def test_no_gradient_doubling(tiny_bert, tiny_data):
    cand, his = _batch(tiny_data)
    m = _model(tiny_bert, tiny_data)
    loss, _, cv, hv = m(cand, his, None)
    m.flat.zero_grad()
    loss.backward()
    before = m.flat.grad.clone()
    m.collect({}, process_user_grad(m.user_encoder.named_parameters()))    # aliases the live grads upstream
    torch.testing.assert_close(m.flat.grad, before)


def test_dedup_off_gives_same_gradients(tiny_bert, tiny_data):
    cand, his = _batch(tiny_data)
    a = _model(tiny_bert, tiny_data)
    b = _model(tiny_bert, tiny_data, dedup=False)
    b.load_state_dict(a.state_dict())
    la = TrainStep(a, impl="torch").run(cand, his)
    lb = TrainStep(b, impl="torch").run(cand, his)
    torch.testing.assert_close(la, lb, atol=1e-6, rtol=1e-5)
    torch.testing.assert_close(a.flat.grad, b.flat.grad, atol=2e-6, rtol=1e-4)


def test_adam_matches_torch_optim(tiny_bert, tiny_data):
    cand, his = _batch(tiny_data)
    a = _model(tiny_bert, tiny_data)
    ref = copy.deepcopy({n: p.detach().clone() for n, p in a.named_parameters() if p.requires_grad})
    params = [torch.nn.Parameter(v.clone()) for v in ref.values()]
    opt = torch.optim.Adam(params, lr=a.cfg.lr)
    step = TrainStep(a, impl="torch")
    for _ in range(3):
        step.run(cand, his)
        for p, n in zip(params, ref.keys()):
            p.grad = a.flat.view(n, grad=True).clone()
        opt.step()
        a.adam_step()
    for p, n in zip(params, ref.keys()):
        torch.testing.assert_close(a.flat.view(n), p.data, atol=1e-7, rtol=1e-5)
        assert not torch.equal(p.data, ref[n])


def test_training_moves_metrics(tiny_bert, tiny_data):
    """Planted-signal data: a few hundred steps must beat chance AUC clearly."""
    cfg = FedRecConfig(dropout=0.0, lr=2e-3)
    torch.manual_seed(0)
    m = UserModel(None, tiny_data.news_index, "cpu", config=cfg, bert_config=tiny_bert)
    tab = ImpressionTable(tiny_data.train_sam, tiny_data.nid2index)
    bs = BatchSampler(tab, 16, seed=0, pin=False)
    tr = Trainer(m, [], snapshot_path="")
    first = None
    for i in range(120):
        cand, his = bs.next()
        loss = tr.step.run(cand, his)
        m.adam_step()
        first = first if first is not None else float(loss)
    v = tr.validate(tiny_data.train_sam, tiny_data.nid2index)
    assert float(loss) < first - 0.05
    assert v[1] > 0.7, v            # AUC on the training impressions
    # upstream's return value: mean loss + the LAST sample's metrics rounded to 2 decimals (client.py:171)
    last = tr.validate(tiny_data.train_sam, tiny_data.nid2index, last_only=True)
    one = tr.validate(tiny_data.train_sam[-1:], tiny_data.nid2index)
    assert abs(last[0] - v[0]) < 1e-6
    assert last[1:] == tuple(round(x, 2) for x in one[1:])


def test_train_on_step_api(tiny_bert, tiny_data):
    m = _model(tiny_bert, tiny_data)
    from fedrec_b200.data import TrainDataset
    from torch.utils.data import DataLoader
    ds = TrainDataset(tiny_data.train_sam[:12], tiny_data.nid2index, tiny_data.news_index, seed=0)
    dl = DataLoader(ds, batch_size=4)
    before = m.flat.param.clone()
    total = train_on_step(m, dl, None)
    assert np.isfinite(float(total)) and m.cnt == 3
    assert not torch.equal(before, m.flat.param)
