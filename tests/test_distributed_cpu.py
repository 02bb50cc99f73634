"""gloo, world_size=2, CPU: the aggregation plumbing of BASELINE.json config #1."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, mode, tmp, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    torch.set_num_threads(2)
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.engine.federated import FederatedRunner
    from fedrec_b200.models.distilbert import DistilBertConfigLite
    bert = DistilBertConfigLite(n_layers=1, dim=32, hidden_dim=64, n_heads=2, hidden_size=32)
    cfg = FedRecConfig(mode=mode, synthetic=True, num_news=128, num_train=32, num_valid=4, title_len=8,
                       batch_size=4, total_epochs=2, global_epochs=2, local_steps=2, dropout=0.0, device="cpu",
                       backend="gloo", snapshot_path=os.path.join(tmp, f"snap_{mode}.pt"), log_dir="",
                       server_rank=1, lr=1e-2)
    r = FederatedRunner(cfg, bert_config=bert)
    res = {}
    if mode == "grad_avg":
        loss = r.local_step()
        local = r.model.flat.grad.clone()
        gathered = [torch.zeros_like(local) for _ in range(world)]
        dist.all_gather(gathered, local)
        r.grad_agg.allreduce_mean_()
        res["mean_ok"] = bool(torch.allclose(r.model.flat.grad, sum(gathered) / world, atol=1e-7))
        res["differs"] = bool(not torch.allclose(gathered[0], gathered[1]))
        r.model.adam_step()
        p = r.model.flat.param.clone()
        ps = [torch.zeros_like(p) for _ in range(world)]
        dist.all_gather(ps, p)
        res["params_equal"] = bool(torch.equal(ps[0], ps[1]))
        res["loss"] = float(loss)
        r.close()
    q.put((rank, res))


def _run_mode(rank, world, port, mode, tmp, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    torch.set_num_threads(2)
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.engine.federated import FederatedRunner
    from fedrec_b200.models.distilbert import DistilBertConfigLite
    bert = DistilBertConfigLite(n_layers=1, dim=32, hidden_dim=64, n_heads=2, hidden_size=32)
    cfg = FedRecConfig(mode=mode, synthetic=True, num_news=128, num_train=32, num_valid=4, title_len=8,
                       batch_size=4, total_epochs=2, global_epochs=2, local_steps=2, dropout=0.0, device="cpu",
                       backend="gloo", snapshot_path=os.path.join(tmp, f"snap_{mode}.pt"), log_dir="",
                       server_rank=1, lr=1e-2, client_weights=[3.0, 1.0] if mode == "param_avg" else None)
    r = FederatedRunner(cfg, bert_config=bert)
    init = r.model.flat.param.clone()
    if mode == "param_avg":
        # one manual round to check the weighted mean in closed form
        for _ in range(2):
            r.local_step()
            r.model.adam_step()
        p = r.model.flat.param.clone()
        ps = [torch.zeros_like(p) for _ in range(world)]
        dist.all_gather(ps, p)
        r.param_agg.average_(r.model.flat.param, r.fedavg_weights())
        ok = torch.allclose(r.model.flat.param, 0.75 * ps[0] + 0.25 * ps[1], atol=1e-6)
        q.put((rank, {"weighted_ok": bool(ok), "moved": bool(not torch.equal(init, p))}))
        r.close()
        return
    loss = None
    if mode == "client_server":
        try:
            loss = r.run_client_server()
            p = r.model.flat.param.clone()
            ps = [torch.zeros_like(p) for _ in range(world)]
            dist.all_gather(ps, p)
            q.put((rank, {"equal": bool(torch.equal(ps[0], ps[1])), "moved": bool(not torch.equal(init, p)),
                          "is_server": r.is_server, "loss": loss}))
        finally:
            r.close()


@pytest.mark.parametrize("mode", ["grad_avg"])
def test_gradient_averaging_two_ranks(tmp_path, mode):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, mode, str(tmp_path), q)) for r in range(2)]
    [p.start() for p in procs]
    out = dict(q.get(timeout=240) for _ in range(2))
    [p.join(60) for p in procs]
    for r in range(2):
        assert out[r]["mean_ok"] and out[r]["params_equal"], out
    assert out[0]["differs"]                              # the clients really had different gradients


def _run_tcp(rank, world, port, tcp_port, tmp, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    torch.set_num_threads(2)
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.engine.federated import FederatedRunner
    from fedrec_b200.models.distilbert import DistilBertConfigLite
    bert = DistilBertConfigLite(n_layers=1, dim=32, hidden_dim=64, n_heads=2, hidden_size=32)
    cfg = FedRecConfig(mode="client_server", synthetic=True, num_news=128, num_train=32, num_valid=4, title_len=8,
                       batch_size=4, total_epochs=1, global_epochs=2, dropout=0.0, device="cpu", backend="gloo",
                       snapshot_path="", log_dir=os.path.join(tmp, f"r{rank}"), server_rank=2, lr=1e-2,
                       transport="tcp", server_host="127.0.0.1", server_port=tcp_port)
    r = FederatedRunner(cfg, bert_config=bert)
    init = r.model.flat.param.clone()
    try:
        r.run_client_server()
        q.put((rank, {"moved": bool(not torch.equal(init, r.model.flat.param)), "is_server": r.is_server,
                      "files": sorted(os.listdir(cfg.log_dir)) if r.is_server else []}))
    finally:
        r.close()


def test_client_server_over_tcp_three_ranks(tmp_path):
    """Upstream's transport: clients upload state_dicts over a socket, the server averages the files (rank 2 = server)."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port, tcp_port = _free_port(), _free_port()
    procs = [ctx.Process(target=_run_tcp, args=(r, 3, port, tcp_port, str(tmp_path), q)) for r in range(3)]
    [p.start() for p in procs]
    out = dict(q.get(timeout=300) for _ in range(3))
    [p.join(60) for p in procs]
    assert out[2]["is_server"] and not out[0]["is_server"]
    assert out[2]["moved"]                                            # the server ended up with the clients' average
    assert "received_model_1.pt" in out[2]["files"] and "received_model_2.pt" in out[2]["files"]


@pytest.mark.parametrize("mode", ["param_avg", "client_server"])
def test_parameter_modes_two_ranks(tmp_path, mode):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_run_mode, args=(r, 2, port, mode, str(tmp_path), q)) for r in range(2)]
    [p.start() for p in procs]
    out = dict(q.get(timeout=240) for _ in range(2))
    [p.join(60) for p in procs]
    if mode == "param_avg":
        assert all(out[r]["weighted_ok"] and out[r]["moved"] for r in range(2)), out
    else:
        assert out[0]["equal"] and out[1]["equal"] and out[1]["is_server"] and not out[0]["is_server"], out
        assert out[0]["moved"] and out[1]["moved"]        # the server received the clients' average


def _run_faulty(rank, world, port, tmp, fault, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port), FEDREC_FAULT=fault)
    torch.set_num_threads(2)
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.engine.federated import FederatedRunner
    from fedrec_b200.models.distilbert import DistilBertConfigLite
    bert = DistilBertConfigLite(n_layers=1, dim=32, hidden_dim=64, n_heads=2, hidden_size=32)
    cfg = FedRecConfig(mode="grad_avg", synthetic=True, num_news=128, num_train=32, num_valid=4, title_len=8, batch_size=4,
                       total_epochs=4, save_every=1, dropout=0.0, device="cpu", backend="gloo", timeout_s=20.0,
                       snapshot_path=os.path.join(tmp, "snap.pt"), log_dir="")
    try:
        r = FederatedRunner(cfg, bert_config=bert)
        start = r.epochs_run
        r.run()
        q.put((rank, {"ok": True, "start": start}))
    except BaseException as e:  # noqa: BLE001
        q.put((rank, {"ok": False, "err": type(e).__name__}))


def test_client_failure_is_detected_and_training_resumes_from_snapshot(tmp_path):
    """A client dies at round 2 (fault injection): the survivor gets an error instead of hanging, and a relaunch resumes
    from the last snapshot instead of starting over (upstream: 'if a client fails the entire training will fail')."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_run_faulty, args=(r, 2, port, str(tmp_path), "kill:1:2", q)) for r in range(2)]
    [p.start() for p in procs]
    [p.join(120) for p in procs]
    assert procs[1].exitcode == 17                                   # the injected fault
    got = {}
    while not q.empty():
        r, res = q.get()
        got[r] = res
    assert 0 in got and not got[0]["ok"], got                         # rank 0 saw an error within the timeout, not a hang
    snap = torch.load(os.path.join(str(tmp_path), "snap.pt"), weights_only=False)
    assert snap["EPOCHS_RUN"] == 2                                    # rounds 0 and 1 were saved
    q2 = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_run_faulty, args=(r, 2, port, str(tmp_path), "", q2)) for r in range(2)]
    [p.start() for p in procs]
    out = dict(q2.get(timeout=240) for _ in range(2))
    [p.join(60) for p in procs]
    assert out[0]["ok"] and out[1]["ok"] and out[0]["start"] == 2 and out[1]["start"] == 2
    assert torch.load(os.path.join(str(tmp_path), "snap.pt"), weights_only=False)["EPOCHS_RUN"] == 4


def _store_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    import time
    from fedrec_b200.parallel import init_distributed, shutdown_distributed
    from fedrec_b200.parallel.comm import agree_all, agree_max, host_barrier
    info = init_distributed("gloo", 30.0, "cpu")
    res = {}
    try:
        t0 = time.monotonic()
        if rank == 1:
            time.sleep(0.5)                              # a late rank costs the others host time, nothing else
        host_barrier("t", 10.0)
        res["waited"] = time.monotonic() - t0
        res["all_true"] = agree_all(True, "a", 10.0)
        res["one_false"] = agree_all(rank != 1, "b", 10.0)
        res["max"] = agree_max(10 + 3 * rank, "m", 10.0)
        host_barrier("t", 10.0)                          # tags are reusable: every call is its own generation
        if rank == 0:
            try:
                host_barrier("only-rank-0", 0.3)
                res["timeout"] = "no error"
            except TimeoutError as e:
                res["timeout"] = str(e)
    except Exception as e:  # noqa: BLE001
        res["error"] = repr(e)
    q.put((rank, res))
    shutdown_distributed(info)


def test_store_barrier_agreement_and_timeout_naming_the_missing_rank():
    """The host-side alignment the aggregation kernels rely on (no GPU work, no collective kernel): barrier, AND, MAX over
    the job's key-value store, and a time-out that names who did not arrive (upstream's only tool is a two-day collective
    timeout, client.py:227)."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_store_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    out = dict(q.get(timeout=120) for _ in range(2))
    [p.join(30) for p in procs]
    for r in range(2):
        assert "error" not in out[r], out[r]
        assert out[r]["all_true"] is True and out[r]["one_false"] is False and out[r]["max"] == 13
    assert out[0]["waited"] >= 0.45
    assert "rank(s) [1]" in out[0]["timeout"], out[0]["timeout"]


def test_torchrun_restarts_the_job_after_a_client_died_and_it_resumes(tmp_path):
    """Elastic recovery the torchrun way: ``--max-restarts 1``; rank 1 dies in epoch 2, the agent relaunches both ranks,
    they pick up the snapshot and finish epochs 2 and 3 (upstream follows the same load-if-present pattern,
    main.py:113-115, but every rank writes the file and nothing ever injects a fault)."""
    import subprocess
    env = dict(os.environ, FEDREC_FAULT="kill:1:2")
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "TORCHELASTIC_RESTART_COUNT"):
        env.pop(k, None)
    for attempt in range(2):                     # one retry: the launcher's own rendezvous is outside this repo's control
        work = os.path.join(str(tmp_path), f"try{attempt}")
        os.makedirs(work)
        snap = os.path.join(work, "snap.pt")
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--max-restarts", "1",
               "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "Gradient_Averaging_main.py"),
               "4", "4", "1", "--synthetic", "--num_news", "120", "--num_train", "16", "--num_valid", "2", "--title_len", "8",
               "--device", "cpu", "--backend", "gloo", "--bert_layers", "1", "--bert_dim", "64", "--bert_heads", "4",
               "--bert_ffn", "128", "--snapshot_path", snap, "--log_dir", "", "--timeout_s", "60", "--heartbeat_s", "0"]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=420, env=env, cwd=work)
        out = r.stdout + r.stderr
        if r.returncode == 0:
            break
    assert r.returncode == 0, out[-3000:]
    assert out.count("fault injection: rank 1 exits at round 2") == 1
    assert "Resuming training from snapshot at Epoch 2" in out
    assert "epoch=3" in out
    assert torch.load(snap, weights_only=False)["EPOCHS_RUN"] == 4


def _run_partial(rank, world, port, mode, q):
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    torch.set_num_threads(2)
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.engine.federated import FederatedRunner
    from fedrec_b200.models.distilbert import DistilBertConfigLite
    bert = DistilBertConfigLite(n_layers=1, dim=32, hidden_dim=64, n_heads=2, hidden_size=32)
    cfg = FedRecConfig(mode=mode, synthetic=True, num_news=128, num_train=16, num_valid=4, title_len=8, batch_size=4,
                       total_epochs=2 if mode == "param_avg" else 1, global_epochs=5, steps_per_epoch=3, local_steps=2,
                       dropout=0.0, device="cpu", backend="gloo", snapshot_path="", log_dir="", lr=1e-2,
                       participation=0.5, server_rank=world - 1, heartbeat_s=0.0)
    r = FederatedRunner(cfg, bert_config=bert)
    n_rounds = 6 if mode == "param_avg" else 5
    sets = [r.active_clients(k) for k in range(n_rounds)]
    r.run_parameter_averaging() if mode == "param_avg" else r.run_client_server()
    p = r.model.flat.param.clone()
    ps = [torch.zeros_like(p) for _ in range(world)]
    dist.all_gather(ps, p)
    res = {"sets": sets, "steps": int(r.model.flat.step), "equal": all(torch.equal(ps[0], x) for x in ps)}
    r.close()
    q.put((rank, res))


@pytest.mark.parametrize("mode", ["param_avg", "client_server"])
def test_client_sampling_trains_a_fraction_of_the_clients_per_round(mode):
    """``participation`` (FedAvg's client fraction): every rank draws the same active set per round, only those clients
    train, everybody ends the round with the same model."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    world = 3 if mode == "param_avg" else 4                     # client_server: 3 clients + the server (last rank)
    procs = [ctx.Process(target=_run_partial, args=(r, world, port, mode, q)) for r in range(world)]
    [p.start() for p in procs]
    out = dict(q.get(timeout=300) for _ in range(world))
    [p.join(60) for p in procs]
    sets = out[0]["sets"]
    assert all(out[r]["sets"] == sets for r in range(world))                  # no communication needed to agree
    assert all(len(s) == 2 and all(c < 3 for c in s) for s in sets) and len({tuple(s) for s in sets}) > 1
    per_round = 2 if mode == "param_avg" else 3 * 1                          # local_steps | steps_per_epoch x total_epochs
    for r in range(3):
        assert out[r]["steps"] == per_round * sum(r in s for s in sets), (r, out[r]["steps"], sets)
    if mode == "client_server":
        assert out[3]["steps"] == 0
    if mode == "param_avg":
        assert all(out[r]["equal"] for r in range(world))
