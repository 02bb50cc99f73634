"""Real multi-GPU checks (one process per GPU, NCCL for plumbing, NVLink kernels for aggregation).

Skipped on boxes with fewer than two GPUs; the single-GPU emulation in test_gpu_aggregate.py covers the kernels
there. Oracle: NCCL all-reduce of the same buffers."""
import os
import socket
import sys

import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    import torch.distributed as dist
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.engine.federated import FederatedRunner
    from fedrec_b200.data import make_synthetic_mind
    res = {}
    try:
        cfg = FedRecConfig(mode="grad_avg", synthetic=True, num_news=2000, num_train=64, num_valid=8, title_len=30,
                           batch_size=8, total_epochs=2, dropout=0.0, clip_norm=2.0, noise_multiplier=0.0,
                           snapshot_path="", log_dir="", transport="nvlink")
        r = FederatedRunner(cfg)
        assert r.symm is not None and r.grad_agg.transport == "nvlink"
        r.local_step()
        g = r.model.flat.grad
        mine = g.clone()
        sc = min(1.0, 2.0 / float(mine.norm()))
        want = mine * sc
        dist.all_reduce(want)
        want /= world
        r.grad_agg.allreduce_mean_()
        torch.cuda.synchronize()
        res["grad_rel_err"] = float((g - want).norm() / want.norm())
        r.model.adam_step()
        # parameter averaging with weights + broadcast
        p = r.model.flat.param
        p.add_(torch.randn_like(p) * 0.01 * (rank + 1))
        w = [float(i + 1) for i in range(world)]
        wantp = p.clone() * (w[rank] / sum(w))
        dist.all_reduce(wantp)
        r.param_agg.average_(p, w)
        torch.cuda.synchronize()
        res["param_rel_err"] = float((p - wantp).norm() / wantp.norm())
        gathered = [torch.zeros_like(p) for _ in range(world)]
        dist.all_gather(gathered, p)
        res["identical"] = all(torch.equal(gathered[0], x) for x in gathered)
        loss = r.run_gradient_averaging()                   # aggregation + Adam fused in one kernel per round
        res["loss"] = loss
        p2 = r.model.flat.param
        g2 = [torch.zeros_like(p2) for _ in range(world)]
        dist.all_gather(g2, p2)
        res["identical_after_training"] = all(torch.equal(g2[0], x) for x in g2)
        r.close()
    except Exception as e:  # noqa: BLE001
        import traceback
        res["error"] = f"{e}\n{traceback.format_exc()}"
    q.put((rank, res))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_nvlink_aggregation_matches_nccl():
    world = min(torch.cuda.device_count(), 8)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    [p.start() for p in procs]
    out = dict(q.get(timeout=600) for _ in range(world))
    [p.join(60) for p in procs]
    for r in range(world):
        assert "error" not in out[r], out[r].get("error")
        assert out[r]["grad_rel_err"] < 1e-5, out
        assert out[r]["param_rel_err"] < 1e-5, out
        assert out[r]["identical"] and out[r]["identical_after_training"]
        assert out[r]["loss"] == out[r]["loss"]


def _worker_frozen(rank, world, port, q, backend):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port), FEDREC_SYMM=backend)
    import torch.distributed as dist
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.engine.federated import FederatedRunner
    res = {}
    try:
        cfg = FedRecConfig(mode="param_avg", synthetic=True, num_news=500, num_train=32, num_valid=4, title_len=30,
                           batch_size=4, total_epochs=1, local_steps=1, dropout=0.0, snapshot_path="", log_dir="",
                           transport="nvlink", average_frozen=True)
        r = FederatedRunner(cfg)
        res["backend"] = r.symm.backend
        res["nvls"] = bool(r.symm.mc_ptr)
        ff = r.model.flat_frozen.param
        assert ff.numel() >= 66_362_880
        ff.add_(0.001 * (rank + 1))
        want = ff.clone()
        dist.all_reduce(want)
        want /= world
        r._average_frozen(r.fedavg_weights())
        torch.cuda.synchronize()
        res["frozen_rel_err"] = float((ff - want).norm() / want.norm())
        # the module still sees the averaged values through its views, and the packed bf16 copies were refreshed
        w = r.model.text_encoder.DistillBert.transformer.layer[0].ffn.lin1.weight
        res["view_ok"] = bool(w.data_ptr() == r.model.flat_frozen.view(
            "text_encoder.DistillBert.transformer.layer.0.ffn.lin1.weight").data_ptr())
        res["loss"] = r.run_parameter_averaging()
        r.close()
    except Exception as e:  # noqa: BLE001
        import traceback
        res["error"] = f"{e}\n{traceback.format_exc()}"
    q.put((rank, res))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("backend", ["ipc", "vmm"])
def test_full_state_dict_average_over_nvlink(backend):
    world = min(torch.cuda.device_count(), 8)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_frozen, args=(r, world, port, q, backend)) for r in range(world)]
    [p.start() for p in procs]
    out = dict(q.get(timeout=600) for _ in range(world))
    [p.join(60) for p in procs]
    for r in range(world):
        assert "error" not in out[r], out[r].get("error")
        assert out[r]["backend"] == backend
        assert out[r]["frozen_rel_err"] < 1e-6, out
        assert out[r]["view_ok"] and out[r]["loss"] == out[r]["loss"]
    print("nvls multicast in use:", out[0]["nvls"])
