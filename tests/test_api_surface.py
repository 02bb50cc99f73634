"""Upstream-named modules and symbols import and behave (SURVEY.md §2.1 inventory)."""
import os
import subprocess
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_upstream_module_names_resolve():
    import attention, dataset, encoder, evaluation_functions, model  # noqa: E401
    assert model.UserModel and model.ServerUserModel
    assert encoder.TextEncoder and encoder.UserEncoder
    assert attention.AdditiveAttention and attention.ScaledDotProductAttention and attention.MultiHeadAttention
    for n in ("TrainDataset", "NewsDataset", "NewsPartDataset", "NewsUpdatorDataset", "UserDataset", "newsample"):
        assert hasattr(dataset, n)
    assert dataset.npratio == 4 and dataset.max_his_len == 50
    for n in ("dcg_score", "ndcg_score", "mrr_score", "compute_amn", "evaluation_split", "roc_auc_score"):
        assert hasattr(evaluation_functions, n)


def test_entrypoint_modules_export_training_loop_names():
    import client, server  # noqa: E401
    import main as main_mod
    import Gradient_Averaging_main as ga
    import Parameter_Averaging_main as pa
    for m in (main_mod, ga, pa, client):
        for n in ("train_on_step", "Trainer", "process_news_grad", "process_user_grad", "main"):
            assert hasattr(m, n), (m.__name__, n)
    assert client.send_saved_model and server.handle_client and server.average_all_params_rcvd


def test_server_model_constructs_without_data(tiny_bert):
    from model import ServerUserModel
    m = ServerUserModel("cpu", bert_config=tiny_bert)                 # upstream raises TypeError here
    assert m.news_index is None and len(list(m.parameters())) > 0


def test_get_and_gen_news_vecs(tiny_bert, tiny_data):
    from model import UserModel
    m = UserModel(None, tiny_data.news_index, "cpu", bert_config=tiny_bert)
    ids = np.array([[3, 4, 3], [7, 3, 0]])
    v = m.gen_news_vecs(ids)
    assert v.shape == (2, 3, 400)
    torch.testing.assert_close(v[0, 0], v[1, 1])                      # same id, same vector (de-dup is exact)
    torch.testing.assert_close(m.get_news_vecs([3])[0], v[0, 0])


def test_synthetic_userdata_cli_and_loader(tmp_path):
    out = str(tmp_path / "UserData")
    r = subprocess.run([sys.executable, "-m", "fedrec_b200.data.synthetic", "--out", out, "--num_news", "200",
                        "--num_train", "16", "--num_valid", "4", "--title_len", "10"], cwd=ROOT, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    from fedrec_b200.data import load_user_data
    nid2index, news_index, train_sam, valid_sam = load_user_data(out)
    assert news_index.shape == (200, 2, 10) and len(train_sam) == 16 and len(valid_sam) == 4 and nid2index["<unk>"] == 0


def test_single_process_entrypoint_runs_and_resumes(tmp_path):
    env = dict(os.environ, PYTHONPATH=ROOT)
    common = ["--synthetic", "--num_news", "128", "--num_train", "32", "--num_valid", "4", "--title_len", "8", "--bert_layers", "1",
              "--bert_dim", "32", "--bert_ffn", "64", "--bert_heads", "2", "--device", "cpu", "--log_dir", str(tmp_path / "logs"),
              "--snapshot_path", str(tmp_path / "snapshot.pt")]
    r = subprocess.run([sys.executable, os.path.join(ROOT, "main.py"), "2", "4", "1"] + common, env=env, cwd=str(tmp_path),
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    snap = torch.load(tmp_path / "snapshot.pt", weights_only=False)
    assert snap["EPOCHS_RUN"] == 2 and "MODEL_STATE" in snap
    r = subprocess.run([sys.executable, os.path.join(ROOT, "main.py"), "3", "4", "1"] + common, env=env, cwd=str(tmp_path),
                       capture_output=True, text=True)
    assert r.returncode == 0 and "Resuming training from snapshot at Epoch 2" in r.stdout
    assert r.stdout.count("training_loss") == 1                      # only epoch 2 ran: no repeated epoch


def test_precomputed_vector_evaluation_path(tiny_bert, tiny_data):
    """encode_all_news + encode_users + evaluation_split == scoring through the model directly."""
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.models import UserModel
    from fedrec_b200.utils.metrics import evaluation_split
    torch.manual_seed(0)
    m = UserModel(None, tiny_data.news_index, "cpu", config=FedRecConfig(dropout=0.0), bert_config=tiny_bert).eval()
    nv = m.encode_all_news(batch=64)
    assert nv.shape == (300, 400)
    torch.testing.assert_close(nv[[5, 7]], m.gen_news_vecs([5, 7]), atol=1e-5, rtol=1e-4)
    uv = m.encode_users(tiny_data.valid_sam, tiny_data.nid2index, nv)
    assert uv.shape == (len(tiny_data.valid_sam), 400)
    samples = [(s[0], [s[1]], s[2], s[3], s[4]) for s in tiny_data.valid_sam]
    res = evaluation_split(nv.numpy(), uv.numpy(), samples, tiny_data.nid2index)
    assert res.shape == (len(samples), 4) and np.isfinite(res).all()
    # same user vector as the training-path forward
    s0 = tiny_data.valid_sam[0]
    his = [tiny_data.nid2index[n] for n in s0[3]][-50:]
    his = his + [0] * (50 - len(his))
    cand = [tiny_data.nid2index[s0[1]]] + [0] * 4
    score, _, hv = m(np.array([cand]), np.array([his]), None, compute_loss=False)
    torch.testing.assert_close(m.user_encoder(hv)[0], uv[0], atol=1e-5, rtol=1e-4)


def test_evaluate_script_runs(tmp_path):
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "evaluate.py"), "--synthetic", "--num_news", "128", "--num_train", "8",
                        "--num_valid", "6", "--title_len", "8", "--bert_layers", "1", "--bert_dim", "32", "--bert_ffn", "64",
                        "--bert_heads", "2", "--device", "cpu", "--snapshot_path", ""], env=env, cwd=str(tmp_path),
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-1500:]
    assert "auc=" in r.stdout and "ndcg@10=" in r.stdout


def test_device_batched_evaluation_equals_the_per_impression_loop(tiny_bert, tiny_data):
    """evaluate.py ranks padded batches with a validity mask; upstream's loop (evaluation_functions.py:33-46) ranks one
    impression at a time on the host. Same numbers."""
    import importlib.util
    import numpy as np
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.models import UserModel
    from fedrec_b200.utils.metrics import evaluation_split
    spec = importlib.util.spec_from_file_location("_evaluate_script", os.path.join(ROOT, "evaluate.py"))
    ev = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ev)
    torch.manual_seed(0)
    m = UserModel(None, tiny_data.news_index, "cpu", config=FedRecConfig(title_len=12, device="cpu"), bert_config=tiny_bert)
    sam = list(tiny_data.valid_sam) + list(tiny_data.train_sam[:20])
    sam = [(s[0], [s[1]] if not isinstance(s[1], (list, tuple)) else list(s[1]), list(s[2])[: 3 + i % 9], s[3], s[4])
           for i, s in enumerate(sam)]
    sam.append((999, ["N100001"], [], [], "u"))                    # single-class impression: skipped by both
    nv = m.encode_all_news()
    uv = m.encode_users(sam, tiny_data.nid2index, nv)
    a = ev.ranked_on_device(nv, uv, sam, tiny_data.nid2index, batch=7)
    b = evaluation_split(nv.numpy(), uv.numpy(), sam, tiny_data.nid2index)
    assert a.shape == b.shape == (len(sam) - 1, 4)
    np.testing.assert_allclose(a, b, atol=2e-5)


def test_upstream_toy_shard_loads_and_trains(tiny_bert):
    """The four files upstream ships in UserData/ (T=50 titles, 76-click history) load and drive a step unchanged."""
    import pytest
    d = "/root/reference/UserData"
    if not os.path.isdir(d):
        pytest.skip("upstream toy shard not available")
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.data import BatchSampler, ImpressionTable, TrainDataset, load_user_data
    from fedrec_b200.engine import TrainStep
    from fedrec_b200.models import UserModel
    nid2index, news_index, train_sam, valid_sam = load_user_data(d)
    assert news_index.shape == (225, 2, 50) and len(train_sam) == 4 and len(valid_sam) == 1
    ds = TrainDataset(train_sam, nid2index, news_index, seed=0)
    cand, his, label = ds[0]
    assert cand.shape == (5,) and his.shape == (50,) and label == 0      # 76-click history truncated to the last 50
    assert list(his) == [nid2index[n] for n in train_sam[0][3]][-50:]
    m = UserModel(None, news_index, "cpu", config=FedRecConfig(title_len=50, dropout=0.0), bert_config=tiny_bert)
    c, h = BatchSampler(ImpressionTable(train_sam, nid2index), 4, seed=0, pin=False).next()
    loss = TrainStep(m, impl="torch").run(c, h)
    assert torch.isfinite(loss) and float(m.flat.grad.abs().sum()) > 0


def test_shipped_userdata_directory_works_like_upstream(tmp_path):
    """`main.py <epochs> <batch> <save_every>` with no other argument reads ./UserData, as upstream does."""
    import shutil
    shutil.copytree(os.path.join(ROOT, "UserData"), tmp_path / "UserData")
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "main.py"), "1", "4", "1", "--bert_layers", "1", "--bert_dim", "32",
                        "--bert_ffn", "64", "--bert_heads", "2", "--device", "cpu", "--log_dir", ""], env=env, cwd=str(tmp_path),
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-1500:]
    assert (tmp_path / "snapshot.pt").exists() and "training_loss" in r.stdout
