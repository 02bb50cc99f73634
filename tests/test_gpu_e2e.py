"""End to end on planted-signal data: the fused sm_100a path reaches the same ranking quality as the fp32 oracle."""
import pytest
import torch

from fedrec_b200 import ops
from fedrec_b200.config import FedRecConfig
from fedrec_b200.data import BatchSampler, ImpressionTable, make_synthetic_mind
from fedrec_b200.engine import Trainer
from fedrec_b200.models import UserModel
from fedrec_b200.models.distilbert import DistilBertConfigLite

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _train(impl_fused: bool, data, steps=150):
    cfg = FedRecConfig(title_len=30, dropout=0.2, batch_size=32, lr=1e-3, fused=impl_fused, cuda_graph=impl_fused, seed=11)
    torch.manual_seed(0)
    bert = DistilBertConfigLite(n_layers=2)            # full width (768, 12 heads): the kernels are built for it
    m = UserModel(None, data.news_index, DEV, config=cfg, bert_config=bert)
    tr = Trainer(m, [], snapshot_path="")
    assert tr.step.impl == ("fused" if impl_fused else "torch")
    bs = BatchSampler(ImpressionTable(data.train_sam, data.nid2index), 32, seed=5, pin=True)
    dc = torch.zeros(32, 5, dtype=torch.int32, device=DEV)
    dh = torch.zeros(32, 50, dtype=torch.int32, device=DEV)
    first = last = None
    for _ in range(steps):
        c, h = bs.next()
        dc.copy_(c, non_blocking=True); dh.copy_(h, non_blocking=True)
        loss = tr.step.run(dc, dh, train=True)
        m.adam_step()
        first = float(loss) if first is None else first
        last = float(loss)
    return first, last, tr.validate(data.valid_sam, data.nid2index), tr.validate(data.train_sam[:128], data.nid2index)


def test_fused_training_matches_oracle_quality():
    ops.require_native()
    data = make_synthetic_mind(num_news=2000, num_train=512, num_valid=128, title_len=30, seed=21)
    f0, f1, fv, ft = _train(True, data)
    o0, o1, ov, ot = _train(False, data)
    assert f1 < f0 - 0.03 and o1 < o0 - 0.03, (f0, f1, o0, o1)        # both learn
    assert ft[1] > 0.65 and ot[1] > 0.65, (ft, ot)                    # AUC on training impressions well above chance
    for a, b in zip(fv[1:], ov[1:]):                                  # validation AUC, MRR, nDCG@5, nDCG@10
        assert abs(a - b) < 0.06, (fv, ov)
