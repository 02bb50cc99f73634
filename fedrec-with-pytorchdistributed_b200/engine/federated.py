"""Federated orchestration: the five upstream entrypoints as modes of one round engine.

Upstream: ``main.py`` (DDP baseline), ``Gradient_Averaging_main.py``
(DDP + a 4-byte loss all-reduce, :146), ``Parameter_Averaging_main.py`` (local step, then
116 per-tensor all-reduces, :144-148), ``client.py`` / ``server.py`` (control-flag
broadcast, per-tensor model broadcast, local epochs, state_dict upload over TCP, server
mean; client.py:256-292, server.py:72-105).

Here a *round* is: every client draws a batch of ``batch_size`` impressions from its own
shard, runs ``local_steps`` fused steps, and then one aggregation chain runs (in the gradient modes inside the next
step's CUDA graph, beside the frozen trunk):

``ddp`` / ``grad_avg``  gradients of all 16 trainable tensors are averaged (optionally
                        clipped and noised per client) **before** Adam;
``param_avg``           Adam runs locally for ``local_steps`` steps, then parameters are
                        FedAvg-averaged (uniform by default = upstream) and every rank
                        receives the result;
``client_server``       rank ``server_rank`` holds no data; each round it publishes the
                        continue/stop flag and the global model, clients train locally
                        with optional LDP noise, and the "upload + mean + redistribute"
                        is the same FedAvg kernel with server weight 0.
"""
from __future__ import annotations

import os
import time
from typing import List, Optional

import torch
import torch.distributed as dist

from ..config import FedRecConfig
from ..data import BatchSampler, DevicePrefetcher, ImpressionTable, load_user_data, make_synthetic_mind
from ..models import ServerUserModel, UserModel
from ..parallel import GradAggregator, ParamAggregator, init_distributed, shutdown_distributed
from ..utils import checkpoint as ckpt
from ..utils.health import FaultInjector, Heartbeat
from ..utils.logging import MetricsLogger
from ..utils.privacy import get_noise_multiplier
from ..utils.trace import Tracer
from .step import TrainStep
from .trainer import validate_impressions


class FederatedRunner:
    def __init__(self, cfg: FedRecConfig, bert_config=None, news_index=None, nid2index=None,
                 train_sam=None, valid_sam=None):
        self.cfg = cfg
        self.info = init_distributed(cfg.backend, cfg.timeout_s, cfg.device)
        self.rank, self.W, self.device = self.info.rank, self.info.world_size, self.info.device
        self.is_server = cfg.mode == "client_server" and self.W > 1 and self.rank == cfg.server_rank
        torch.manual_seed(cfg.seed)                       # identical initial weights on every rank
        # ---- data: each rank is one client with its own shard (README.md:22) ----------
        if news_index is None:
            # every rank is a client with its own shard: "UserData_{rank}" picks a per-rank directory on a shared box
            cfg.data_dir = cfg.data_dir.format(rank=self.rank, local_rank=self.info.local_rank)
            if cfg.synthetic or not os.path.isdir(cfg.data_dir):
                d = make_synthetic_mind(cfg.num_news, cfg.num_train, cfg.num_valid, cfg.title_len,
                                        cfg.max_his_len, seed=cfg.seed, client=self.rank)
                nid2index, news_index, train_sam, valid_sam = d.nid2index, d.news_index, d.train_sam, d.valid_sam
            else:
                nid2index, news_index, train_sam, valid_sam = load_user_data(cfg.data_dir)
        self.nid2index, self.news_index = nid2index, news_index
        self.train_sam, self.valid_sam = train_sam, valid_sam
        # ---- peer-mapped memory for the aggregation kernels ----------------------------------
        self.symm = None
        allocator = None
        if self.device.type == "cuda" and self.W > 1 and cfg.transport == "nvlink":
            from ..parallel.symm import SymmetricHeap
            heap_mb = 64 + (288 if cfg.average_frozen else 0)       # the full fp32 state_dict is 257.6 MiB
            ok = torch.zeros(1, device=self.device)
            try:
                self.symm = SymmetricHeap.create(self.info, heap_bytes=heap_mb << 20, timeout_s=cfg.agg_timeout_s)
                ok.fill_(1.0)
            except Exception as e:  # noqa: BLE001  (no peer access / IPC not permitted in this container)
                print(f"[fedrec r{self.rank}] peer-mapped heap unavailable ({e}); aggregation falls back to NCCL", flush=True)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)               # every rank must agree on the transport
            if float(ok) < 1.0:
                self.symm = None
            else:
                allocator = self.symm.alloc_f32
        # ---- model -------------------------------------------------------------------
        if self.is_server:
            self.model = ServerUserModel(self.device, config=cfg, bert_config=bert_config, flat_allocator=allocator)
        else:
            self.model = UserModel(None, news_index, self.device, config=cfg, bert_config=bert_config,
                                   flat_allocator=allocator)
        if self.symm is not None:
            try:
                self.symm.finalize()
            except RuntimeError as e:                           # raised on every rank together (see finalize)
                print(f"[fedrec r{self.rank}] {e}; aggregation falls back to NCCL", flush=True)
                self.symm = None
        self.step = None if self.is_server else TrainStep(self.model, cfg)
        transport = "nvlink" if self.symm is not None else "collective"
        participants = [r for r in range(self.W) if not (cfg.mode == "client_server" and self.W > 1 and r == cfg.server_rank)]
        symm = self.symm
        if self.W == 1 and self.device.type == "cuda" and cfg.transport == "nvlink" and cfg.fused:
            from ..parallel.symm import SymmetricHeap
            symm, transport = SymmetricHeap.single(self.device), "local"     # same kernel chain, no peers
            self._local_pad = symm
        self.grad_agg = GradAggregator(self.model.flat, self.W, self.rank, transport, cfg.clip_norm,
                                       cfg.noise_multiplier, cfg.seed, symm, participants=participants,
                                       blocks=cfg.agg_blocks)
        # rounds with long, uneven local phases (a server idling through the clients' epochs, snapshot writes) line the
        # ranks up on the host before the kernel that waits for peers
        self.param_agg = ParamAggregator(self.W, self.rank, "nvlink" if self.symm is not None else "collective",
                                         self.symm, host_align=cfg.mode in ("client_server", "param_avg"),
                                         timeout_s=cfg.timeout_s)
        self.local_opt = None
        self.sampler = None
        self._native = None
        if not self.is_server and train_sam:
            table = ImpressionTable(train_sam, nid2index, cfg.max_his_len)
            self.sampler = BatchSampler(table, cfg.batch_size, cfg.npratio, seed=cfg.seed + 101 * self.rank,
                                        pin=self.device.type == "cuda")
            self.prefetch = DevicePrefetcher(self.sampler, self.device)
            if cfg.native_loader:
                # batches come from the C++ producer thread (csrc/host_loader.cu) in the step's packed layout and enter
                # the step through its host-mapped ring (TrainStep.run_host): no sampler work, staging copy or prefetch
                # stream on the launch path
                from ..data.native_loader import NativeBatchLoader
                self._native = NativeBatchLoader(table, cfg.batch_size, cfg.npratio, seed=cfg.seed + 101 * self.rank)
                self.step.enable_host_io(cfg.batch_size, 1 + cfg.npratio, cfg.max_his_len)
        self.logger = MetricsLogger(cfg.log_dir, cfg.run_name, self.rank, enabled=bool(cfg.log_dir))
        self.tracer = Tracer(enabled=os.environ.get("FEDREC_TRACE") == "1", device=self.device)
        self.faults = FaultInjector()
        self.heartbeat: Optional[Heartbeat] = None
        self.degraded: Optional[str] = None
        if cfg.heartbeat_s > 0 and self.W > 1:
            store = dist.distributed_c10d._get_default_store()
            self.heartbeat = Heartbeat(store, self.rank, self.W, period_s=cfg.heartbeat_s,
                                       stale_s=max(10 * cfg.heartbeat_s, 30.0), on_stale=self._on_stale).start()
        self._tcp_server = None
        self.round = 0
        self.epochs_run = 0
        if cfg.snapshot_path and os.path.exists(cfg.snapshot_path):
            info = ckpt.load_snapshot(cfg.snapshot_path, self.model)
            self.epochs_run = info["EPOCHS_RUN"]
            self.round = int(info.get("ROUND", self.epochs_run))
            print(f"Resuming training from snapshot at Epoch {self.epochs_run}")
        if self.W > 1:
            self._sync_initial_state()
        self._attach_optimizer()

    # ---- optimiser placement ---------------------------------------------------------------------------
    def _attach_optimizer(self) -> None:
        """Pipeline the update into the step where the kernels allow it (fused step + device chain): gradient modes
        use the cross-client chain, the local-step modes a one-rank chain (Adam only). Otherwise the loops below call
        aggregation / Adam after each step, like upstream's ``model.update()`` (main.py:90)."""
        cfg = self.cfg
        self.pipelined = False
        if self.step is None or not cfg.pipeline_optimizer:
            return
        if cfg.mode in ("ddp", "grad_avg"):
            self.pipelined = self.step.attach_optimizer(self.grad_agg, cfg.lr, cfg.beta1, cfg.beta2, cfg.adam_eps)
        elif self.step.impl == "fused":
            from ..parallel.symm import SymmetricHeap
            self.local_opt = GradAggregator(self.model.flat, 1, 0, "local", 0.0, 0.0, cfg.seed,
                                            SymmetricHeap.single(self.device), blocks=cfg.agg_blocks)
            self.pipelined = self.step.attach_optimizer(self.local_opt, cfg.lr, cfg.beta1, cfg.beta2, cfg.adam_eps)

    def flush(self) -> None:
        """Apply the update the pipelined step still owes and surface any aggregation failure."""
        if self.step is not None:
            self.step.flush()
        if self.symm is not None:
            self.symm.check()

    def _local_adam(self) -> None:
        if not self.pipelined:
            self.model.adam_step()

    def _degrade(self, err: Exception, epoch: int) -> int:
        """A peer-memory aggregation gave up (a rank was late beyond ``agg_timeout_s``, or died). Every rank ends up
        here: the abort word stops all chains. Agree on where to resume, put the pads back, continue over NCCL, and
        re-synchronise weights and optimiser state from rank 0. Upstream: "if a client fails the entire training will
        fail" (report p.4); here a late client costs the steps in flight."""
        from ..parallel.comm import agree_max
        print(f"[fedrec r{self.rank}] {err}; continuing with transport=collective", flush=True)
        self.degraded = str(err)
        resume = agree_max(epoch, "degrade", self.cfg.timeout_s)
        if self.step is not None:
            fused = self.step._fused
            if fused is not None:
                fused._pending = False
                fused._opt = None
                fused._graphs.clear()
        self.pipelined = False
        self.symm.reset()
        self.grad_agg.degrade()
        self.param_agg.degrade()
        fl = self.model.flat
        for t in (fl.param, fl.exp_avg, fl.exp_avg_sq):
            dist.broadcast(t, src=0)
        st = torch.tensor([fl.step], device=self.device, dtype=torch.int64)
        dist.all_reduce(st, op=dist.ReduceOp.MAX)
        fl.step = int(st.item())
        self.logger.log({"event": "transport_degraded", "reason": str(err), "resume_epoch": resume}, step=epoch)
        return resume

    # ------------------------------------------------------------------------------------------
    def _on_stale(self, ranks: List[int]) -> None:
        print(f"[fedrec r{self.rank}] no progress from rank(s) {ranks} for >{self.heartbeat.stale_s:.0f}s — "
              f"a client has failed or is a straggler; the job will time out after {self.cfg.timeout_s:.0f}s "
              f"and can be resumed from {self.cfg.snapshot_path}", flush=True)

    def _sync_initial_state(self) -> None:
        """Everyone starts from rank 0's weights (DDP constructor broadcast)."""
        src = self.cfg.server_rank if self.cfg.mode == "client_server" else 0
        src = src if src < self.W else 0
        self.param_agg.broadcast_(self.model.flat.param, src)
        # the frozen trunk is seeded identically on every rank; broadcast only when asked to
        if self.cfg.average_frozen:
            self.param_agg.broadcast_(self.model.flat_frozen.param, src)
            self.model.on_weights_changed()

    def next_batch(self):
        return self.prefetch.next()

    def local_step(self) -> torch.Tensor:
        if self._native is not None:
            with self.tracer.span("step"):
                ticket = self.step.run_host(self._native.next_packed())
            return _HostLoss(self.step, ticket, self.device)
        cand, his = self.next_batch()
        with self.tracer.span("step"):
            return self.step.run(cand, his, train=True)

    def fedavg_weights(self) -> Optional[List[float]]:
        cfg = self.cfg
        if cfg.client_weights is not None:
            w = list(cfg.client_weights)
            if len(w) != self.W:
                raise ValueError(f"client_weights has {len(w)} entries for world size {self.W}")
        else:
            w = [1.0] * self.W
        if cfg.mode == "client_server" and self.W > 1:
            w[cfg.server_rank] = 0.0
        return w

    def active_clients(self, rnd: int) -> List[int]:
        """The clients that train in aggregation round ``rnd`` (FedAvg's client fraction; ``cfg.participation`` < 1). The
        draw is a function of (seed, round), so every rank computes the same set without talking. Everybody still takes
        part in the aggregation itself: the others enter with weight 0 and receive the result."""
        cfg = self.cfg
        clients = [r for r in range(self.W) if not (cfg.mode == "client_server" and self.W > 1 and r == cfg.server_rank)]
        if cfg.participation >= 1.0 or len(clients) <= 1:
            return clients
        import numpy as np
        m = max(1, int(round(cfg.participation * len(clients))))
        g = np.random.default_rng([int(cfg.seed) & 0xFFFFFFFF, 7919, int(rnd)])
        return sorted(int(x) for x in g.choice(clients, size=m, replace=False))

    @staticmethod
    def _round_weights(weights: List[float], active: List[int]) -> List[float]:
        w = [x if r in active else 0.0 for r, x in enumerate(weights)]
        return w if sum(w) > 0 else [1.0 if r in active else 0.0 for r in range(len(weights))]

    def validate(self):
        if self.is_server or not self.valid_sam:
            return None
        self.flush()
        return validate_impressions(self.step, self.model, self.valid_sam, self.nid2index)

    def save(self, epoch: int) -> None:
        self.flush()
        ckpt.save_snapshot(self.cfg.snapshot_path, self.model, next_epoch=epoch + 1, rank=self.rank,
                           extra={"ROUND": self.round})

    # ---- modes ---------------------------------------------------------------------------------------
    def steps_per_epoch(self) -> int:
        """Steps in one epoch: ``cfg.steps_per_epoch`` if set, else one pass over this client's shard
        (ceil(impressions / batch_size); upstream's epoch is a pass over its DataLoader, main.py:55-91,135-139)."""
        if self.cfg.steps_per_epoch > 0:
            return self.cfg.steps_per_epoch
        n = len(self.sampler.table) if self.sampler is not None else 0
        return max(1, -(-n // max(self.cfg.batch_size, 1)))

    def run_gradient_averaging(self) -> float:
        from ..parallel.symm import AggregationError
        cfg = self.cfg
        loss = torch.zeros((), device=self.device)
        spe = self._agreed_steps_per_epoch()
        epoch = self.epochs_run
        while epoch < cfg.total_epochs:
            try:
                self.faults.maybe_fire(self.rank, epoch)
                self._epoch_started(spe)
                for _ in range(spe):
                    loss = self.local_step()                  # pipelined: also applies the previous step's update
                    if not self.pipelined:
                        with self.tracer.span("aggregate+adam"):  # clip, noise, mean over NVLink and the Adam step
                            self.grad_agg.allreduce_mean_and_adam_(cfg.lr, cfg.beta1, cfg.beta2, cfg.adam_eps)
                    self.round += 1
                    if self.heartbeat:
                        self.heartbeat.beat(self.round)
                self._end_of_epoch(epoch, loss, reduce_loss=cfg.mode == "grad_avg")
                epoch += 1
            except AggregationError as e:
                epoch = self._degrade(e, epoch)
        self.flush()
        return float(loss)

    def _agreed_steps_per_epoch(self) -> int:
        """Gradient modes step in lockstep: every client takes as many steps per epoch as the largest shard needs
        (smaller shards wrap around)."""
        spe = self.steps_per_epoch()
        if self.W > 1:
            t = torch.tensor([spe], device=self.device, dtype=torch.int64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            spe = int(t.item())
        return spe

    def run_parameter_averaging(self) -> float:
        from ..parallel.symm import AggregationError
        cfg = self.cfg
        loss = torch.zeros((), device=self.device)
        weights = self.fedavg_weights()
        spe = self._agreed_steps_per_epoch()
        epoch = self.epochs_run
        while epoch < cfg.total_epochs:
            try:
                self.faults.maybe_fire(self.rank, epoch)
                self._epoch_started(spe * max(cfg.local_steps, 1))
                for it in range(spe):
                    active = self.active_clients(self.round)
                    if self.rank in active:
                        for _ in range(max(cfg.local_steps, 1)):
                            loss = self.local_step()
                            self._local_adam()
                    self.flush()
                    if it == spe - 1 and (epoch + 1) % cfg.save_every == 0 and cfg.snapshot_path:
                        self.save(epoch)                              # saved before averaging, as PA:140-141
                    with self.tracer.span("aggregate"):
                        w_round = self._round_weights(weights, active)
                        self.param_agg.average_(self.model.flat.param, w_round)
                        if cfg.average_frozen and self.W > 1:
                            self._average_frozen(w_round)
                    self.round += 1
                    if self.heartbeat:
                        self.heartbeat.beat(self.round)
                self._end_of_epoch(epoch, loss, reduce_loss=False, save=False)
                epoch += 1
            except AggregationError as e:
                epoch = self._degrade(e, epoch)
        self.flush()
        return float(loss)

    def _average_frozen(self, weights) -> None:
        """The frozen trunk too (upstream averages all 116 tensors): one kernel over the 265 MB flat buffer."""
        self.param_agg.average_(self.model.flat_frozen.param, weights)
        self.model.on_weights_changed()

    def run_client_server(self) -> float:
        """Both roles; ``global_epochs`` rounds of ``total_epochs`` local steps (client.py:258-292, server.py:72-105)."""
        from ..parallel.symm import AggregationError
        cfg = self.cfg
        flag = torch.zeros(1, device=self.device)
        weights = self.fedavg_weights()
        src = cfg.server_rank if self.W > 1 else 0
        sigma = None
        if cfg.dp_epsilon > 0 and not self.is_server:
            sample_rate = min(1.0, cfg.batch_size / max(len(self.train_sam), 1))
            sigma = get_noise_multiplier(cfg.dp_epsilon, cfg.dp_delta, sample_rate, epochs=cfg.dp_epochs)
            cfg.news_grad_noise = sigma                          # client.py:87-89: noise on news-vector grads
            print(f"Applying sigma {sigma:.4f} for DP")
        loss = torch.zeros((), device=self.device)
        rnd = self.epochs_run
        while True:
            if self.rank == src:
                flag.fill_(1.0 if rnd < cfg.global_epochs else 0.0)
            if self.W > 1:
                dist.broadcast(flag, src=src)                     # X7 control flag
            elif rnd >= cfg.global_epochs:
                flag.zero_()
            if int(flag.item()) == 0:
                break
            try:
                self.faults.maybe_fire(self.rank, rnd)
                self.param_agg.broadcast_(self.model.flat.param, src)  # X8 model sync
                active = self.active_clients(rnd)
                if not self.is_server and self.rank in active:
                    for _ in range(max(cfg.total_epochs, 1) * self.steps_per_epoch()):   # client.py:284: local epochs per round
                        loss = self.local_step()
                        self._local_adam()
                    self.flush()
                    v = self.validate()
                    rec = {"round": rnd, "training_loss": float(loss)}
                    if v:
                        rec.update({"validation_loss": v[0], "valid_auc": v[1], "valid_mrr": v[2],
                                    "val_ndcg@5": v[3], "val_ndcg@10": v[4]})
                    self.logger.log(rec, step=rnd)
                with self.tracer.span("aggregate"):
                    if cfg.transport == "tcp" and self.W > 1:
                        self._tcp_upload_and_average(src, self._round_weights(weights, active))   # X9 + K23 the upstream way (other hosts)
                    else:
                        self.param_agg.average_(self.model.flat.param, self._round_weights(weights, active))   # X9 + K23 in one kernel chain
                        if self.symm is not None and not self.degraded:
                            self.symm.check(sync=True)            # rounds are long: find out now, not a round later
            except AggregationError as e:
                rnd = self._degrade(e, rnd)                       # every rank lands here; the round is repeated over NCCL
                continue
            rnd += 1
            self.round = rnd
            if self.heartbeat:
                self.heartbeat.beat(rnd)
            if cfg.snapshot_path and rnd % cfg.save_every == 0 and self.rank == src:
                ckpt.save_snapshot(cfg.snapshot_path, self.model, next_epoch=rnd, rank=0, extra={"ROUND": rnd})
        return float(loss)

    def _tcp_upload_and_average(self, src: int, weights) -> None:
        """Clients stream their state_dict to the server over TCP; the server averages the files and loads the result
        (client.py:191-210, 286-290; server.py:25-55, 80-103). The next round's model sync distributes it."""
        from ..utils import transport
        cfg = self.cfg
        out_dir = cfg.log_dir or "."
        if self.rank == src:
            if self._tcp_server is None:
                self._tcp_server = transport.StateDictServer("0.0.0.0", cfg.server_port, out_dir=out_dir)
            dist.barrier()                                              # the server is listening
            paths = self._tcp_server.receive_round(self.W - 1, timeout_s=cfg.timeout_s)
            ids = self._tcp_server.last_client_ids                    # slot (1-based) → uploading rank
            w = [weights[ids.get(i + 1, 0)] for i in range(len(paths))]
            self.model.load_state_dict(transport.average_state_dicts(paths, w if len(set(w)) > 1 else None))
        else:
            dist.barrier()
            transport.send_state_dict(ckpt.model_state_fp32_cpu(self.model), cfg.server_host, cfg.server_port,
                                      client_id=self.rank, timeout_s=cfg.timeout_s)

    def _epoch_started(self, steps: int) -> None:
        self._epoch_t0, self._epoch_steps = time.perf_counter(), steps

    def _end_of_epoch(self, epoch: int, loss: torch.Tensor, reduce_loss: bool, save: bool = True) -> None:
        cfg = self.cfg
        rec = {"epoch": epoch, "training_loss": float(loss)}        # float(): the device has finished the epoch
        t0 = getattr(self, "_epoch_t0", None)
        if t0 is not None:                                          # upstream shows tqdm's it/s (main.py:57)
            dt = max(time.perf_counter() - t0, 1e-9)
            rec["epoch_time_s"] = round(dt, 4)
            rec["impressions_per_s"] = round(self._epoch_steps * cfg.batch_size * (1 if self.is_server else self.W) / dt, 1)
            self._epoch_t0 = None
        if cfg.clip_norm > 0 and cfg.noise_multiplier > 0 and cfg.mode in ("ddp", "grad_avg") and self.train_sam \
                and self.model.flat.step > 0:
            # what the clipped + noised client gradients have cost so far (RDP accountant, sampled Gaussian mechanism
            # with q = batch / shard size); upstream only ever computes a sigma for a target epsilon (client.py:217-226)
            from ..utils.privacy import compute_epsilon
            q = min(1.0, cfg.batch_size / max(len(self.train_sam), 1))
            try:
                rec["dp_epsilon_spent"] = round(compute_epsilon(cfg.noise_multiplier, q, int(self.model.flat.step), cfg.dp_delta), 4)
                rec["dp_delta"] = cfg.dp_delta
            except (OverflowError, ValueError, ZeroDivisionError):
                pass                                                # sigma too small for the accountant's orders: no figure
        if cfg.validate_every > 0 and (epoch + 1) % cfg.validate_every == 0:
            v = self.validate()      # every rank over its own validation shard, like the client role (client.py:180)
            if v:
                rec.update({"validation_loss": v[0], "valid_auc": v[1], "valid_mrr": v[2], "val_ndcg@5": v[3],
                            "val_ndcg@10": v[4]})
        if reduce_loss and self.W > 1:
            tot = loss.detach().clone()
            dist.all_reduce(tot)                                   # GA:146
            rec["loss_sum_all_clients"] = float(tot)
        self.logger.log(rec, step=epoch)
        if save and cfg.snapshot_path and (epoch + 1) % cfg.save_every == 0:
            self.save(epoch)
            if self.W > 1 and self.symm is not None and not self.degraded:
                from ..parallel.comm import host_barrier
                host_barrier("after-save", cfg.timeout_s)          # rank 0 wrote 270 MB: do not let the peers' kernels wait it out

    def run(self) -> float:
        mode = self.cfg.mode
        try:
            if mode in ("ddp", "grad_avg"):
                return self.run_gradient_averaging()
            if mode == "param_avg":
                return self.run_parameter_averaging()
            if mode == "client_server":
                return self.run_client_server()
            raise ValueError(f"unknown mode {mode!r}")
        finally:
            self.close()

    def close(self) -> None:
        try:
            if self.step is not None:
                self.step.flush()
        except Exception:  # noqa: BLE001  (closing after a failure must not raise again)
            pass
        if self.heartbeat:
            self.heartbeat.stop()
        if self._native is not None:
            self._native.close()
            self._native = None
        if self.tracer.enabled:
            os.makedirs(self.cfg.log_dir or ".", exist_ok=True)
            self.tracer.export_chrome(os.path.join(self.cfg.log_dir or ".", f"trace.rank{self.rank}.json"))
        self.logger.close()
        if self.symm is not None:
            self.symm.close()
        if self._tcp_server is not None:
            self._tcp_server.close()
        shutdown_distributed(self.info)


class _HostLoss:
    """Loss of a step that ran through ``TrainStep.run_host``: read from host memory when (and if) somebody asks."""

    def __init__(self, step, ticket: int, device):
        self._step, self._ticket, self._device, self._value = step, ticket, device, None

    def __float__(self) -> float:
        if self._value is None:
            self._value = float(self._step.loss_of(self._ticket))
        return self._value

    def item(self) -> float:
        return float(self)

    def detach(self) -> torch.Tensor:
        return torch.tensor(float(self), device=self._device)


_POSITIONAL = {
    "ddp": ["total_epochs", "batch_size", "save_every"],
    "grad_avg": ["total_epochs", "batch_size", "save_every"],
    "param_avg": ["total_epochs", "batch_size", "save_every"],
    "client": ["total_epochs", "batch_size", "save_every", "dp_epsilon", "run_name"],
    "server": ["global_epochs"],
}


def run_entrypoint(kind: str, argv: List[str]) -> float:
    """Shared ``__main__`` of the five reference-named scripts."""
    mode = {"client": "client_server", "server": "client_server"}.get(kind, kind)
    cfg = FedRecConfig.from_argv(argv, _POSITIONAL[kind], mode=mode)
    if kind == "server":
        rank = int(os.environ.get("RANK", "0"))
        world = int(os.environ.get("WORLD_SIZE", "1"))
        if world > 1 and rank != cfg.server_rank:
            raise SystemExit(f"server.py must run as rank {cfg.server_rank} (got {rank}); clients run client.py")
    return FederatedRunner(cfg).run()
