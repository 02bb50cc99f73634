"""Per-round aggregation: gradient averaging and parameter averaging.

What is being replaced: DDP's bucketed gradient all-reduce (X3), the
4-byte loss all-reduce (X5), the 116 per-tensor parameter all-reduces of
Parameter_Averaging_main.py:144-148 (X6), and the client→server state_dict upload +
server mean + per-tensor broadcast of client.py/server.py (X7-X9, K23).

Transports behind one interface:

``collective``  torch.distributed (gloo on CPU, NCCL on GPU) over the flat buffer:
                one all-reduce instead of 16 or 116. This is the *baseline* path and
                the one CPU tests use.
``nvlink``      the sm_100a kernels in ``csrc/aggregate.cu``: each rank's flat buffer
                lives in peer-mapped memory; a short chain of kernels does L2-clip,
                Gaussian noise, the cross-GPU reduction over NVLink (peer loads or
                NVLS multimem), the 1/N (or FedAvg weight) scaling, the
                write-back/broadcast and Adam. No NCCL call.
``local``       one rank on a GPU: the same chain without the cross-GPU part.

The ``nvlink`` and ``local`` chains can be **pipelined** into the training step
(``pipelined`` property; ``engine/fused.py``): the step marks its gradients as pending
and the chain that consumes them runs at the start of the *next* step on a side
branch, next to the frozen DistilBERT trunk, which depends on nothing the aggregation
writes. Upstream's analogue is DDP overlapping its buckets with backward (main.py:69-70).

Semantics adopted: every trainable gradient is averaged
before the optimiser step; clipping bounds each client's flat gradient to L2 norm C;
noise is N(0, (sigma·C)²) per coordinate per client, added before averaging.
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch
import torch.distributed as dist

from .. import ops


class GradAggregator:
    def __init__(self, flat, world_size: int, rank: int, transport: str = "collective",
                 clip_norm: float = 0.0, noise_multiplier: float = 0.0, seed: int = 0, symm=None,
                 group=None, participants: Optional[Sequence[int]] = None, blocks: int = 0):
        self.flat = flat
        self.W, self.rank = world_size, rank
        self.transport = transport
        self.clip, self.sigma = float(clip_norm), float(noise_multiplier)
        self.seed = seed
        self.symm = symm
        self.group = group
        self.round = 0
        self.blocks = blocks
        self.participants = list(participants) if participants is not None else list(range(world_size))
        if transport in ("nvlink", "local") and symm is not None and flat is not None:
            symm.set_step(flat.step)

    @property
    def noise_std(self) -> float:
        # sigma is relative to the clip bound when clipping is on (DP-SGD convention)
        return self.sigma * (self.clip if self.clip > 0 else 1.0)

    @property
    def on_device_chain(self) -> bool:
        return self.transport in ("nvlink", "local") and self.symm is not None

    def _zero_if_not_participating(self, g: torch.Tensor) -> None:
        # the peer-memory kernels sum every rank's buffer: a rank outside `participants` (the data-less server)
        # must contribute zeros, exactly like the collective path below
        if len(self.participants) != self.W and self.rank not in self.participants:
            g.zero_()

    def chain_(self, lr: float, b1: float, b2: float, eps: float, gated: bool = False) -> None:
        """Enqueue clip + noise + mean + Adam on the current stream (device chains only). ``gated`` chains take the
        Adam step count from the device (``symm.set_step``) and only run when the step marked gradients as pending."""
        fl = self.flat
        assert self.on_device_chain
        self.symm.check()
        self._zero_if_not_participating(fl.grad)
        self.round += 1
        fl.step += 1
        self.symm.grad_allreduce_adam_(fl.grad, clip=self.clip, noise_std=self.noise_std, seed=self.seed,
                                       scale=1.0 / len(self.participants), params=fl.param, exp_avg=fl.exp_avg,
                                       exp_avg_sq=fl.exp_avg_sq, step=0 if gated else fl.step, lr=lr, b1=b1, b2=b2,
                                       eps=eps, gated=gated, blocks=self.blocks)

    def allreduce_mean_and_adam_(self, lr: float, b1: float, b2: float, eps: float) -> torch.Tensor:
        """Clip + noise + mean over participants + one Adam step on the flat parameters.

        On the device transports this is one chain of kernels without a host round trip; otherwise aggregation
        followed by the separate fused Adam kernel. Either way ``flat.step`` advances by one."""
        from .. import ops as _ops
        fl = self.flat
        if self.on_device_chain and fl.param.is_cuda:
            self.chain_(lr, b1, b2, eps, gated=False)
            return fl.grad
        self.allreduce_mean_()
        _ops.adam_flat(fl, lr, b1, b2, eps)
        return fl.grad

    def allreduce_mean_(self, buf: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Clip + noise this client's gradient, then replace it by the mean over participants."""
        g = self.flat.grad if buf is None else buf
        self.round += 1
        if self.transport == "nvlink" and self.symm is not None and self.W > 1:
            self.symm.check()
            self._zero_if_not_participating(g)
            self.symm.grad_allreduce_(g, clip=self.clip, noise_std=self.noise_std, seed=self.seed,
                                      scale=1.0 / len(self.participants), blocks=self.blocks)
            return g
        ops.clip_noise_(g, self.clip, self.noise_std, seed=self.seed + 7919 * self.round, stream_id=self.rank)
        if self.W > 1:
            if self.rank not in self.participants:
                g.zero_()
            dist.all_reduce(g, op=dist.ReduceOp.SUM, group=self.group)
            g.mul_(1.0 / len(self.participants))
        return g

    def degrade(self) -> None:
        """Switch to the collective transport (after a reported aggregation failure; every rank calls this together)."""
        self.transport = "collective"


class ParamAggregator:
    """Weighted FedAvg of a flat parameter buffer + broadcast of the result to every rank."""

    def __init__(self, world_size: int, rank: int, transport: str = "collective", symm=None, group=None,
                 host_align: bool = False, timeout_s: float = 1800.0):
        self.W, self.rank = world_size, rank
        self.transport = transport
        self.symm = symm
        self.group = group
        self.host_align = host_align          # line ranks up on the host first (rounds with long, uneven local phases)
        self.timeout_s = timeout_s

    def _align(self) -> None:
        if self.host_align and self.W > 1:
            from .comm import host_barrier
            host_barrier("param-agg", self.timeout_s)

    def average_(self, buf: torch.Tensor, weights: Optional[Sequence[float]] = None) -> torch.Tensor:
        """``buf ← Σ_i w_i·buf_i / Σ_i w_i`` on every rank (weight 0 = aggregation-only server rank)."""
        if self.W == 1:
            return buf
        w = [1.0] * self.W if weights is None else [float(x) for x in weights]
        tot = sum(w)
        if tot <= 0:
            raise ValueError("FedAvg weights must have a positive sum")
        mine = w[self.rank] / tot
        if self.transport == "nvlink" and self.symm is not None:
            self.symm.check()
            self._align()
            self.symm.param_average_(buf, [x / tot for x in w])
            return buf
        buf.mul_(mine)
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=self.group)
        return buf

    def broadcast_(self, buf: torch.Tensor, src: int) -> torch.Tensor:
        if self.W > 1:
            if self.transport == "nvlink" and self.symm is not None:
                self.symm.check()
                self._align()
                w = [0.0] * self.W
                w[src] = 1.0
                self.symm.param_average_(buf, w)
            else:
                dist.broadcast(buf, src=src, group=self.group)
        return buf

    def degrade(self) -> None:
        self.transport = "collective"
