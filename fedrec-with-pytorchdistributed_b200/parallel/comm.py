"""Process-group plumbing: one process per GPU (or per CPU client under gloo).

Reference: every entrypoint calls ``init_process_group(backend="gloo")`` (main.py:144,
Gradient_Averaging_main.py:153, Parameter_Averaging_main.py:156, client.py:227,
server.py:59) and never selects a device per rank.  Here the backend
is NCCL when CUDA is present (gloo otherwise), each rank binds ``cuda:LOCAL_RANK``,
and a single-process run needs no launcher at all.
"""
from __future__ import annotations

import datetime
import os
from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist


@dataclass
class DistInfo:
    rank: int
    local_rank: int
    world_size: int
    backend: str
    device: torch.device
    owns_group: bool


def init_distributed(backend: str = "auto", timeout_s: float = 1800.0, device: Optional[str] = None) -> DistInfo:
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    use_cuda = torch.cuda.is_available() and (device is None or str(device).startswith("cuda") or device == "auto")
    if backend == "auto":
        backend = "nccl" if use_cuda else "gloo"
    if use_cuda:
        dev = torch.device("cuda", local_rank % torch.cuda.device_count())
        torch.cuda.set_device(dev)
    else:
        dev = torch.device("cpu")
    owns = False
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29531")
        kwargs = dict(backend=backend, rank=rank, world_size=world,
                      timeout=datetime.timedelta(seconds=timeout_s))
        if backend == "nccl":
            kwargs["device_id"] = dev
        attempt = int(os.environ.get("TORCHELASTIC_RESTART_COUNT", "0") or 0)
        if attempt > 0 and os.environ.get("TORCHELASTIC_USE_AGENT_STORE") == "True":
            # relaunched by `torchrun --max-restarts`: the agent's key-value store outlives the workers, and with a static
            # rendezvous the previous incarnation's transport addresses are still in it under the same keys; a restarted
            # rank that reads one of those connects to a dead process. Give every incarnation its own key space.
            base = dist.TCPStore(os.environ["MASTER_ADDR"], int(os.environ["MASTER_PORT"]), world, False,
                                 datetime.timedelta(seconds=timeout_s))
            kwargs["store"] = dist.PrefixStore(f"fedrec/attempt_{attempt}", base)
        dist.init_process_group(**kwargs)
        owns = True
    return DistInfo(rank, local_rank, world, backend, dev, owns)


_BARRIER_SEQ = {}


def host_barrier(tag: str = "b", timeout_s: float = 1800.0, world_size: Optional[int] = None) -> None:
    """Barrier over the job's key-value store: no GPU work, no collective kernel.

    Used to line ranks up on the *host* before a kernel that waits for its peers (the aggregation chain's wait kernel),
    so an arbitrarily late rank (a server idling through the clients' local epochs, rank 0 writing a snapshot) costs
    host time on the others instead of a GPU-side spin. Upstream's only alignment is the blocking gloo collective
    itself (Parameter_Averaging_main.py:146, client.py:257). Raises ``TimeoutError`` naming the missing ranks."""
    if not dist.is_initialized():
        return
    W = world_size or dist.get_world_size()
    if W <= 1:
        return
    import time
    store = dist.distributed_c10d._get_default_store()
    n = _BARRIER_SEQ.get(tag, 0)
    _BARRIER_SEQ[tag] = n + 1
    key = f"fedrec/hb/{tag}/{n}"
    store.set(f"{key}/r{dist.get_rank()}", b"1")
    arrived = store.add(key, 1)
    t0, sleep = time.monotonic(), 0.0002
    while arrived < W:
        if time.monotonic() - t0 > timeout_s:
            missing = []
            for r in range(W):
                try:
                    if not store.check([f"{key}/r{r}"]):
                        missing.append(r)
                except Exception:  # noqa: BLE001
                    missing.append(r)
            raise TimeoutError(f"host barrier '{tag}' #{n}: rank(s) {missing} did not arrive within {timeout_s:.0f}s")
        time.sleep(sleep)
        sleep = min(sleep * 1.5, 0.01)
        arrived = store.add(key, 0)


def agree_all(ok: bool, tag: str = "agree", timeout_s: float = 1800.0) -> bool:
    """True iff every rank passed ``ok=True`` (store-based AND; doubles as a barrier)."""
    if not dist.is_initialized() or dist.get_world_size() <= 1:
        return ok
    store = dist.distributed_c10d._get_default_store()
    n = _BARRIER_SEQ.get("agree/" + tag, 0)
    _BARRIER_SEQ["agree/" + tag] = n + 1
    bad = store.add(f"fedrec/agree/{tag}/{n}", 0 if ok else 1)
    host_barrier("agree/" + tag + "/b", timeout_s)
    bad = store.add(f"fedrec/agree/{tag}/{n}", 0)
    return bad == 0


def agree_max(value: int, tag: str = "max", timeout_s: float = 1800.0) -> int:
    """Largest ``value`` over all ranks (store-based; doubles as a barrier). Used when ranks notice a failure at
    different points of their host loops and must resume from a common one."""
    if not dist.is_initialized() or dist.get_world_size() <= 1:
        return value
    store = dist.distributed_c10d._get_default_store()
    n = _BARRIER_SEQ.get("amax/" + tag, 0)
    _BARRIER_SEQ["amax/" + tag] = n + 1
    store.set(f"fedrec/amax/{tag}/{n}/r{dist.get_rank()}", str(int(value)))
    host_barrier("amax/" + tag + "/b", timeout_s)
    return max(int(store.get(f"fedrec/amax/{tag}/{n}/r{r}").decode()) for r in range(dist.get_world_size()))


def shutdown_distributed(info: Optional[DistInfo] = None) -> None:
    if dist.is_initialized() and (info is None or info.owns_group):
        dist.destroy_process_group()
