"""Metrics logging: JSON lines on disk, optional wandb when the user opts in through the environment.

The reference prints and calls wandb with a key committed in source (client.py:214-219,
182-189). Nothing of that is kept: no credential lives here, logging is
local by default, and ``WANDB_API_KEY`` + ``FEDREC_WANDB=1`` turn wandb on if it is installed.
"""
from __future__ import annotations

import json
import os
import time
from typing import Any, Dict, Optional


class MetricsLogger:
    def __init__(self, log_dir: str = "logs", run_name: str = "fedrec", rank: int = 0, enabled: bool = True,
                 echo: bool = True):
        self.rank = rank
        self.enabled = enabled
        self.echo = echo and os.environ.get("FEDREC_QUIET") != "1"
        self.path = None
        self._fh = None
        self._wandb = None
        if enabled:
            os.makedirs(log_dir, exist_ok=True)
            self.path = os.path.join(log_dir, f"{run_name}.rank{rank}.jsonl")
            self._fh = open(self.path, "a", buffering=1)
        if enabled and os.environ.get("FEDREC_WANDB") == "1" and os.environ.get("WANDB_API_KEY"):
            try:
                import wandb  # type: ignore
                wandb.init(project=os.environ.get("WANDB_PROJECT", "fedrec_b200"), name=run_name)
                self._wandb = wandb
            except Exception as e:  # pragma: no cover
                print(f"[fedrec] wandb unavailable: {e}")

    def log(self, record: Dict[str, Any], step: Optional[int] = None) -> None:
        rec = {"t": time.time(), "rank": self.rank}
        if step is not None:
            rec["step"] = step
        rec.update({k: (float(v) if hasattr(v, "item") else v) for k, v in record.items()})
        if self._fh:
            self._fh.write(json.dumps(rec) + "\n")
        if self.echo and self.rank == 0:
            body = " ".join(f"{k}={v:.5g}" if isinstance(v, float) else f"{k}={v}" for k, v in rec.items()
                            if k not in ("t", "rank"))
            print(f"[fedrec r{self.rank}] {body}", flush=True)
        if self._wandb is not None:
            self._wandb.log({k: v for k, v in rec.items() if k not in ("t", "rank")}, step=step)

    def close(self) -> None:
        if self._fh:
            self._fh.close()
            self._fh = None
