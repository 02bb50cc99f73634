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
        dist.init_process_group(**kwargs)
        owns = True
    return DistInfo(rank, local_rank, world, backend, dev, owns)


def shutdown_distributed(info: Optional[DistInfo] = None) -> None:
    if dist.is_initialized() and (info is None or info.owns_group):
        dist.destroy_process_group()
