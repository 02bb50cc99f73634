import os

import torch

from fedrec_b200.config import FedRecConfig
from fedrec_b200.models import UserModel
from fedrec_b200.utils import checkpoint as ckpt


def test_snapshot_roundtrip_and_container(tmp_path, tiny_bert, tiny_data):
    cfg = FedRecConfig()
    a = UserModel(None, tiny_data.news_index, "cpu", config=cfg, bert_config=tiny_bert)
    a.flat.grad.normal_()
    a.adam_step()
    path = str(tmp_path / "snapshot.pt")
    assert ckpt.save_snapshot(path, a, next_epoch=4, rank=0, extra={"ROUND": 4})
    assert not ckpt.save_snapshot(str(tmp_path / "other.pt"), a, next_epoch=4, rank=1)      # only rank 0 writes
    assert not os.path.exists(tmp_path / "other.pt")
    raw = torch.load(path, weights_only=False)
    assert {"MODEL_STATE", "EPOCHS_RUN"} <= set(raw) and raw["EPOCHS_RUN"] == 4
    assert list(raw["MODEL_STATE"].keys()) == list(a.state_dict().keys())
    assert all(v.dtype == torch.float32 for v in raw["MODEL_STATE"].values() if v.is_floating_point())
    b = UserModel(None, tiny_data.news_index, "cpu", config=cfg, bert_config=tiny_bert)
    info = ckpt.load_snapshot(path, b)
    assert info["EPOCHS_RUN"] == 4 and info["ROUND"] == 4
    for n in a.flat.names:
        torch.testing.assert_close(a.flat.view(n), b.flat.view(n))
    torch.testing.assert_close(a.flat.exp_avg, b.flat.exp_avg)
    assert b.flat.step == 1
    # parameters are still views of the flat buffer after loading
    for n, p in b.named_parameters():
        if p.requires_grad:
            assert p.data_ptr() == b.flat.view(n).data_ptr()


def test_loads_upstream_style_snapshot(tmp_path, tiny_bert, tiny_data):
    """A file holding only the two upstream keys (main.py:127-133) must load."""
    a = UserModel(None, tiny_data.news_index, "cpu", bert_config=tiny_bert)
    path = str(tmp_path / "ref.pt")
    torch.save({"MODEL_STATE": a.state_dict(), "EPOCHS_RUN": 2}, path)
    b = UserModel(None, tiny_data.news_index, "cpu", bert_config=tiny_bert)
    assert ckpt.load_snapshot(path, b)["EPOCHS_RUN"] == 2
    for n in a.flat.names:
        torch.testing.assert_close(a.flat.view(n), b.flat.view(n))


def test_bare_state_dict_files(tmp_path, tiny_bert, tiny_data):
    a = UserModel(None, tiny_data.news_index, "cpu", bert_config=tiny_bert)
    p = str(tmp_path / "model.pt")
    ckpt.save_state_dict(p, a)
    sd = ckpt.load_state_dict_file(p)
    assert list(sd.keys()) == list(a.state_dict().keys())
