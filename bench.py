#!/usr/bin/env python
"""Headline benchmark: training impressions/sec on N GPUs of one box (BASELINE.json).

Config (BASELINE.json #2 + #4): `Gradient_Averaging_main` semantics, one federated client per GPU,
50 users (impressions) per client per round, history 50, 4 negatives per positive, title length 30,
full-size DistilBERT text encoder + user encoder, per-client L2 clip + LDP Gaussian noise inside the
gradient aggregation, bf16 tensor-core compute, synthetic MIND-shaped data, random-init weights.
One step = one round: forward, backward, gradient aggregation across clients, Adam. Weak scaling.

  python bench.py --gpus 1 --steps 20 --warmup 3
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
         --master-port 29500 bench.py --gpus 8 --steps 20 --warmup 3
  python bench.py --impl reference ...      # the unmodified upstream code from baseline/_ref

Rank 0 prints exactly one JSON line on stdout.
"""
from __future__ import annotations

import argparse
import contextlib
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

B_USERS, HIS_LEN, NPRATIO, TITLE_LEN, NUM_NEWS = 50, 50, 4, 30, 65536
CLIP, SIGMA = 2.0, 0.05


# ---- clocks -----------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi polled in a child process (no thread of ours competes with the launch loop). ``window()`` brackets a
    timed region on the wall clock; the reported median covers only samples that fall inside such windows (all samples
    when none was marked or none fell inside), so idle time between two loops does not flatter the figure."""
    Q = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int, period_ms: int = 25):
        self.path = tempfile.mktemp(prefix="fedrec_clocks_", suffix=".csv")
        self.proc = None
        self.idx = gpu_index
        self.period_ms = period_ms
        self.windows = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", str(self.period_ms), "-i", str(self.idx)], stdout=open(self.path, "w"),
                                         stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None
        return self

    @contextlib.contextmanager
    def window(self):
        t0 = time.time()
        try:
            yield
        finally:
            self.windows.append((t0, time.time()))

    @staticmethod
    def _stamp(text: str):
        import datetime
        try:
            return datetime.datetime.strptime(text, "%Y/%m/%d %H:%M:%S.%f").timestamp()
        except ValueError:
            return None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        rows, mx = [], []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    rows.append((self._stamp(f[0]), float(f[1]), [nm for nm, val in zip(names, f[5:9])
                                                                   if val.lower().startswith("active")]))
                    mx.append(float(f[2]))
                except ValueError:
                    continue
            os.remove(self.path)
        except OSError:
            pass
        inside = [r for r in rows if r[0] is not None and any(a <= r[0] <= b for a, b in self.windows)]
        used = inside or rows
        sm = sorted(r[1] for r in used)
        reasons = sorted({nm for r in used for nm in r[2]})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm), "samples_total": len(rows),
                "scope": "inside the timed regions" if inside else "whole run (no sample fell inside a timed region)"}


class _NoWindow:
    @contextlib.contextmanager
    def window(self):
        yield


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def synthetic_news_index(np, seed=0):
    from fedrec_b200.data import make_synthetic_mind
    return make_synthetic_mind(num_news=NUM_NEWS, num_train=8, num_valid=1, title_len=TITLE_LEN, seed=seed).news_index


def make_batches(np, n_batches, rank, seed=0):
    """Zipf-popular news ids like MIND impressions: [n, B, 1+K] candidates and [n, B, L] histories (int64, host)."""
    rng = np.random.default_rng(seed + 1000 * rank)
    p = 1.0 / np.arange(1, NUM_NEWS) ** 1.05
    p /= p.sum()
    cand = rng.choice(np.arange(1, NUM_NEWS), size=(n_batches, B_USERS, 1 + NPRATIO), p=p)
    his = rng.choice(np.arange(1, NUM_NEWS), size=(n_batches, B_USERS, HIS_LEN), p=p)
    lens = rng.integers(10, HIS_LEN + 1, size=(n_batches, B_USERS))
    his[np.arange(HIS_LEN)[None, None, :] >= lens[:, :, None]] = 0
    return cand.astype(np.int64), his.astype(np.int64)


def finish(args, impl, rank, world, ms_per_step, e2e_ms, h2d, d2h, launches, clocks, extra_cfg):
    import torch
    import torch.distributed as dist
    if world > 1:
        t = torch.tensor([ms_per_step, e2e_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_per_step, e2e_ms = float(t[0]), float(t[1])
    if rank != 0:
        return
    value = world * B_USERS / (ms_per_step * 1e-3)
    e2e_val = world * B_USERS / (e2e_ms * 1e-3)
    # `config` names the workload only and is identical in every arm; what an arm does on top of it goes to `details`
    cfg = {"model": "FedRec joint DistilBERT-base text encoder + user encoder (MHA 20x20 + additive attention)",
           "mode": "Gradient_Averaging_main: per-round gradient averaging",
           "global_batch": world * B_USERS, "users_per_client_per_round": B_USERS, "seq_len": TITLE_LEN,
           "history_len": HIS_LEN, "negatives_per_positive": NPRATIO, "num_news": NUM_NEWS,
           "parallelism": f"dp{world} (one federated client per GPU)",
           "l2_policy": "working set per step (>1 GB activations + 130 MB bf16 weights) exceeds the 126 MB L2; "
                        "a fresh batch of ids every step"}
    details = {"privacy": ("per-client L2 clip C=%g + LDP Gaussian noise sigma=%g inside the aggregation" % (CLIP, SIGMA)
                           if impl != "reference" else "none (upstream as is: no clip, no noise)")}
    details.update(extra_cfg)
    out = {"metric": "training impressions/sec/box", "value": round(value, 2), "unit": "impressions/s",
           "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4),
           "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32" if impl == "reference" else "bf16",
           "data": "synthetic MIND-shaped impressions (Zipf news popularity), random-init weights",
           "impl": impl, "config": cfg, "details": details, "clocks": clocks,
           "e2e": {"value": round(e2e_val, 2), "unit": "impressions/s", "ms_per_step": round(e2e_ms, 4),
                   "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
           "gpu_launches": int(launches)}
    print(json.dumps(out), flush=True)


# ---- our framework -------------------------------------------------------------------------------------
def pack_ids(torch, cand_h, his_h):
    """[n, B*L + B*C] int32, histories first: the step's one input buffer (FusedStep.run_packed)."""
    n = cand_h.shape[0]
    return torch.cat([torch.from_numpy(his_h).reshape(n, -1), torch.from_numpy(cand_h).reshape(n, -1)], 1).to(torch.int32)


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        print(json.dumps({"impl": "ours", "unavailable": "no CUDA device visible"}))
        return
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}; launch with torch.distributed.run"
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.engine.federated import FederatedRunner
    from fedrec_b200.ops import dispatch as D
    cfg = FedRecConfig(mode=args.mode, local_steps=args.local_steps, average_frozen=args.average_frozen, batch_size=B_USERS, max_his_len=HIS_LEN, npratio=NPRATIO, title_len=TITLE_LEN,
                       num_news=NUM_NEWS, clip_norm=CLIP, noise_multiplier=SIGMA, synthetic=True, snapshot_path="",
                       log_dir="", transport=args.transport, num_train=8, num_valid=1, seed=1234,
                       trunk_cache=args.trunk_cache, pipeline_optimizer=args.pipeline == "on",
                       agg_timeout_s=args.agg_timeout_s, heartbeat_s=0.0, cuda_graph=args.graph == "on",
                       ln_fold={"on": True, "off": False}.get(args.ln_fold, FedRecConfig().ln_fold))
    news_index = synthetic_news_index(np)
    runner = FederatedRunner(cfg, news_index=news_index, nid2index={"<unk>": 0}, train_sam=[], valid_sam=[])
    dev = runner.device
    total = args.warmup + args.steps
    cand_h, his_h = make_batches(np, total, rank)
    ids_pin = pack_ids(torch, cand_h, his_h).pin_memory()           # [total, B*(L+C)] int32, pinned
    ids_d = ids_pin.to(dev)
    B, C, L = B_USERS, 1 + NPRATIO, HIS_LEN

    weights = runner.fedavg_weights()
    state = {"n": 0}

    fault = [int(x) for x in args.fault.split(":")] if args.fault else None      # rank:call:seconds, fires once

    def one_step(ids):
        state["calls"] = state.get("calls", 0) + 1
        if fault and rank == fault[0] and state["calls"] == fault[1]:
            time.sleep(fault[2])                        # testing only: this rank goes silent (see --fault)
        loss = runner.step.run_packed(ids, B, C, L)
        if args.mode == "grad_avg":                     # BASELINE.json config 2/4: average gradients, then Adam
            if not runner.pipelined:                    # pipelined: the step itself applies the previous update
                c_ = runner.cfg
                runner.grad_agg.allreduce_mean_and_adam_(c_.lr, c_.beta1, c_.beta2, c_.adam_eps)
        else:                                           # config 3: local Adam steps, FedAvg every `local_steps`
            runner._local_adam()
            state["n"] += 1
            if state["n"] % args.local_steps == 0:
                runner.step.flush()
                runner.param_agg.average_(runner.model.flat.param, weights)
                if args.average_frozen:
                    runner._average_frozen(weights)
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    from collections import deque
    from fedrec_b200.parallel.symm import AggregationError

    def settle():
        """Everything queued has run; an aggregation that gave up on a peer raises here, on every rank, before any rank
        enters a collective the others would never reach."""
        runner.step.flush()
        torch.cuda.synchronize()
        if runner.symm is not None:
            runner.symm.check()

    def guarded(fn):
        try:
            return fn()
        except AggregationError as e:          # a peer was late beyond --agg_timeout_s: finish the measurement over NCCL
            runner._degrade(e, 0)
            return fn()

    # ---- device-timed: ids already on the device ----
    def timed_device():
        for i in range(args.warmup):
            one_step(ids_d[i])
        settle()
        barrier()
        l0 = D.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with win():
            e0.record()
            for i in range(args.warmup, total):
                one_step(ids_d[i])
            runner.step.flush()                         # the last step's update belongs to the timed region
            e1.record()
            settle()
        barrier()
        return e0.elapsed_time(e1) / args.steps, D.launch_count() - l0

    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    win = (sampler or _NoWindow()).window
    # ---- end to end through the public API (TrainStep.run_host / loss_of): every step the ids travel from pinned host
    #      memory to the device (pulled over PCIe by the step's first kernel) and the step's loss travels back into host
    #      memory (pushed by its last kernel) where the host reads it. The host reads the loss of step i while step
    #      i + e2e_lag is already queued, so it never drains the launch queue ----
    ids_host = ids_pin.numpy()
    step = runner.step
    step.enable_host_io(B, C, L, slots=max(4, args.e2e_lag + 2))

    def one_step_host(i):
        t = step.run_host(ids_host[i])
        if args.mode == "grad_avg":
            if not runner.pipelined:
                c_ = runner.cfg
                runner.grad_agg.allreduce_mean_and_adam_(c_.lr, c_.beta1, c_.beta2, c_.adam_eps)
        else:
            runner._local_adam()
            state["n"] += 1
            if state["n"] % args.local_steps == 0:
                runner.step.flush()
                runner.param_agg.average_(runner.model.flat.param, weights)
                if args.average_frozen:
                    runner._average_frozen(weights)
        return t

    def timed_e2e():
        for i in range(args.warmup):                    # untimed warm-up; the first call captures the host-I/O graph
            step.loss_of(one_step_host(i))
        settle()
        barrier()
        tickets = deque()
        t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
        with win():
            t0.record()
            host_loss, n_read = 0.0, 0
            for i in range(args.warmup, total):
                tickets.append(one_step_host(i))
                if len(tickets) > args.e2e_lag:
                    host_loss, n_read = step.loss_of(tickets.popleft()), n_read + 1
            runner.step.flush()
            while tickets:
                host_loss, n_read = step.loss_of(tickets.popleft()), n_read + 1
            t1.record()
            settle()
        barrier()
        assert n_read == args.steps
        return t0.elapsed_time(t1) / args.steps, host_loss

    # the end-to-end loop is the headline and runs first; the device-timed loop (ids resident on the device) follows
    e2e_ms, host_loss = guarded(timed_e2e)
    if args.cooldown_s > 0:                             # let the board's power average recover: both loops then start alike
        time.sleep(args.cooldown_s)
    ms, launches = guarded(timed_device)
    clocks = sampler.stop() if sampler else None
    h2d = ids_pin[0].numel() * 4
    symm = runner.symm
    finish(args, "ours", rank, world, ms, e2e_ms, h2d, 8, launches, clocks,
           {"transport": runner.grad_agg.transport, "step_impl": runner.step.impl, "ln_fold": bool(cfg.ln_fold),
            "optimizer": ("pipelined into the next step's graph (aggregation + Adam beside the frozen trunk)"
                          if runner.pipelined else "after the step"),
            "e2e_loss_read_lag_steps": args.e2e_lag, "last_loss": host_loss,
            "order": "end-to-end loop first, %g s idle, device-timed loop second (under the 1 kW power cap a loop that "
                     "follows another back to back sees lower clocks)" % args.cooldown_s,
            "degraded": runner.degraded,
            "peer_heap": (symm.describe() if symm else "none"),
            **({"trunk_cache": "ON — trunk outputs cached across steps; not comparable with the headline number"}
               if args.trunk_cache else {}),
            **({} if args.mode == "grad_avg" else
               {"mode": f"Parameter_Averaging_main: FedAvg every {args.local_steps} local steps"
                        f"{' incl. the frozen trunk (270 MB)' if args.average_frozen else ' (trainable state)'}"})})
    runner.close()


# ---- strong baseline: this repo's algorithm on library kernels ---------------------------------------------------
def run_torch(args):
    """Same algorithm as ``ours`` (per-step de-dup, one batched frozen-trunk pass, no second trunk pass, clip + noise +
    mean + Adam) on library kernels only: PyTorch modules under bf16 autocast (cuBLAS GEMMs, SDPA attention), NCCL
    all-reduce, torch ops for clip/noise, torch.optim.Adam(fused). Isolates what the hand-written kernels earn."""
    import numpy as np
    import torch
    import torch.distributed as dist
    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        print(json.dumps({"impl": "torch", "unavailable": "no CUDA device visible"}))
        return
    from fedrec_b200.config import FedRecConfig
    from fedrec_b200.engine.step import TrainStep
    from fedrec_b200.models import UserModel
    from fedrec_b200.parallel import init_distributed
    info = init_distributed()
    dev = info.device
    cfg = FedRecConfig(mode="grad_avg", batch_size=B_USERS, max_his_len=HIS_LEN, npratio=NPRATIO, title_len=TITLE_LEN,
                       num_news=NUM_NEWS, clip_norm=CLIP, noise_multiplier=SIGMA, synthetic=True, snapshot_path="",
                       log_dir="", transport="collective", seed=1234, fused=False)
    torch.manual_seed(cfg.seed)
    model = UserModel(None, synthetic_news_index(np), dev, config=cfg)
    for mod in model.text_encoder.DistillBert.modules():
        if hasattr(mod, "use_sdpa"):
            mod.use_sdpa = True
    step = TrainStep(model, cfg, impl="torch")
    fl = model.flat
    opt = torch.optim.Adam([p for p in model.parameters() if p.requires_grad], lr=cfg.lr, fused=True)
    total = args.warmup + args.steps
    cand_h, his_h = make_batches(np, total, rank)
    cand_pin = torch.from_numpy(cand_h).pin_memory(); his_pin = torch.from_numpy(his_h).pin_memory()
    cand_d, his_d = cand_pin.to(dev), his_pin.to(dev)
    std = SIGMA * CLIP

    def one_step(c, h):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = step.run(c, h, train=True)
        g = fl.grad
        g.mul_(torch.clamp(CLIP / g.norm().clamp_min(1e-30), max=1.0))
        g.add_(torch.randn_like(g), alpha=std)
        if world > 1:
            dist.all_reduce(g)
            g.mul_(1.0 / world)
        opt.step()
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(12):                                 # library autotuning and allocator growth settle (not counted as warm-up)
        one_step(cand_d[i % total], his_d[i % total])
    for i in range(args.warmup):
        one_step(cand_d[i], his_d[i])
    barrier()
    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.warmup, total):
        one_step(cand_d[i], his_d[i])
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / args.steps
    barrier()
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    host_loss = 0.0
    for i in range(args.warmup, total):
        host_loss = float(one_step(cand_pin[i].to(dev, non_blocking=True), his_pin[i].to(dev, non_blocking=True)).item())
    t1.record()
    barrier()
    clocks = sampler.stop() if sampler else None
    h2d = (cand_pin[0].numel() + his_pin[0].numel()) * 8
    finish(args, "torch", rank, world, ms, t0.elapsed_time(t1) / args.steps, h2d, 4, 0, clocks,
           {"note": "this repo's algorithm (per-step de-dup, one frozen-trunk pass, head-only backward) on library kernels: "
                    "PyTorch bf16 autocast (cuBLAS, SDPA), NCCL all-reduce, torch clip/noise, torch.optim.Adam(fused)",
            "last_loss": host_loss})
    if dist.is_initialized():
        dist.destroy_process_group()


# ---- the upstream code, unmodified, from baseline/_ref ---------------------------------------------------------
def run_reference(args):
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isfile(os.path.join(ref_dir, "Gradient_Averaging_main.py")):
        print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref is missing (run baseline/install_ref.sh)"}))
        return
    import numpy as np
    import torch
    if not torch.cuda.is_available():
        print(json.dumps({"impl": "reference", "unavailable": "no CUDA device visible"}))
        return
    import torch.distributed as dist
    rank, local_rank, world = dist_env()
    from baseline.harness import load_reference
    torch.cuda.set_device(local_rank)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29533")
    os.environ.setdefault("RANK", "0"); os.environ.setdefault("LOCAL_RANK", "0"); os.environ.setdefault("WORLD_SIZE", "1")
    if not dist.is_initialized():
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
    ga = load_reference(ref_dir)                       # the upstream Gradient_Averaging_main module
    news_index = synthetic_news_index(np)
    dev = torch.device("cuda", local_rank)
    torch.manual_seed(1234)
    with contextlib.redirect_stdout(sys.stderr):
        model = ga.UserModel(ga.NewsDataset(news_index), news_index, dev).to(dev)
        optimizer = ga.optim.SGD(model.parameters(), lr=0.00005)
        trainer = ga.Trainer(model, [], optimizer, save_every=10 ** 9, snapshot_path="/nonexistent/snapshot.pt")
    total = args.warmup + args.steps
    cand_h, his_h = make_batches(np, total, rank)
    label = torch.zeros(B_USERS, dtype=torch.long, device=dev)   # upstream leaves labels on the CPU

    def one_step(i):
        dl = [(torch.from_numpy(cand_h[i]), torch.from_numpy(his_h[i]), label)]
        with contextlib.redirect_stdout(sys.stderr):
            return ga.train_on_step(trainer.model, dl, optimizer)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(args.warmup):
        one_step(i)
    barrier()
    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.warmup, total):
        one_step(i)
    e1.record()
    barrier()
    clocks = sampler.stop() if sampler else None
    ms = e0.elapsed_time(e1) / args.steps
    n_rows = B_USERS * (1 + NPRATIO + HIS_LEN)
    h2d = n_rows * 2 * TITLE_LEN * 8 * 2 + n_rows * 400 * 4 * 2      # token rows (two passes) + vectors and grads back up
    d2h = n_rows * 400 * 4 * 2 + 4                                   # news vectors and their grads + loss
    finish(args, "reference", rank, world, ms, ms, h2d, d2h, 0, clocks,
           {"note": "unmodified upstream sources (baseline/_ref) run through Gradient_Averaging_main.Trainer/"
                    "train_on_step with DDP over NCCL; its public API takes host ids and returns a host loss, so the "
                    "device-timed and end-to-end regions coincide; upstream has no clip/noise in this mode"})
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "torch"],
                    help="ours = the sm_100a framework; reference = the unmodified upstream code; torch = this repo's "
                         "algorithm on library kernels (strong baseline)")
    ap.add_argument("--pipeline", choices=["on", "off"], default="on",
                    help="apply step i's aggregation + Adam inside step i+1's graph, beside the frozen trunk")
    ap.add_argument("--graph", choices=["on", "off"], default="on", help="replay the step from a CUDA graph")
    ap.add_argument("--e2e_lag", type=int, default=1, help="steps between queuing a loss read-back and consuming it")
    ap.add_argument("--agg_timeout_s", type=float, default=30.0)
    ap.add_argument("--cooldown_s", type=float, default=1.0, help="idle time between the end-to-end and the device-timed loop")
    ap.add_argument("--fault", default="", help="testing: RANK:CALL:SECONDS makes one rank sleep before its CALL-th step, to "
                                                "exercise the time-out report and the fall-back to NCCL")
    ap.add_argument("--transport", default="nvlink", choices=["nvlink", "collective"])
    ap.add_argument("--mode", default="grad_avg", choices=["grad_avg", "param_avg"],
                    help="grad_avg = BASELINE.json configs 2/4 (default, the headline); param_avg = config 3")
    ap.add_argument("--local_steps", type=int, default=5)
    ap.add_argument("--average_frozen", action="store_true")
    ap.add_argument("--ln_fold", choices=["default", "on", "off"], default="default",
                    help="LayerNorm folded into the trunk GEMM epilogues (same arithmetic, fewer kernels)")
    ap.add_argument("--trunk_cache", action="store_true",
                    help="NOT the headline: keep trunk outputs of seen news resident in HBM (skips DistilBERT work for repeats)")
    args = ap.parse_args()
    if args.steps is None:
        args.steps = {"ours": 200, "torch": 50}.get(args.impl, 5)
    if args.gpus > 1 and "RANK" not in os.environ:
        # convenience: re-launch under torch.distributed.run
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29517"), __file__] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    {"ours": run_ours, "reference": run_reference, "torch": run_torch}[args.impl](args)


if __name__ == "__main__":
    main()
