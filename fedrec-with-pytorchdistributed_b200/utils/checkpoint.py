"""Snapshot save / resume.

Container kept from the reference (main.py:127-133):
``{"MODEL_STATE": state_dict, "EPOCHS_RUN": int}`` so a reference snapshot loads
here and the other way round.  Fixes adopted from SURVEY.md §9.11: only one rank
writes, the write is atomic (temp file + rename), ``EPOCHS_RUN`` is the *next*
epoch to run, and optimiser moments / RNG / round counter ride along under extra
keys that the reference loader simply ignores.

``save_state_dict`` / ``load_state_dict_file`` handle the bare state_dict files
exchanged in client/server mode (``model.pt``, ``received_model_{i}.pt``;
client.py:286-288, server.py:27,41-42).
"""
from __future__ import annotations

import os
import tempfile
from typing import Any, Dict, Optional

import torch


def _atomic_save(obj: Any, path: str) -> None:
    d = os.path.dirname(os.path.abspath(path))
    os.makedirs(d, exist_ok=True)
    fd, tmp = tempfile.mkstemp(prefix=".tmp_", suffix=".pt", dir=d)
    os.close(fd)
    try:
        torch.save(obj, tmp)
        os.replace(tmp, path)
    finally:
        if os.path.exists(tmp):
            os.remove(tmp)


def model_state_fp32_cpu(model: torch.nn.Module) -> Dict[str, torch.Tensor]:
    """fp32 CPU copy of the state_dict, in registration order (116 keys)."""
    return {k: v.detach().to("cpu", torch.float32).clone() if v.is_floating_point() else v.detach().cpu().clone()
            for k, v in model.state_dict().items()}


def save_snapshot(path: str, model: torch.nn.Module, next_epoch: int, rank: int = 0,
                  extra: Optional[Dict[str, Any]] = None) -> bool:
    if rank != 0:
        return False
    snap: Dict[str, Any] = {"MODEL_STATE": model_state_fp32_cpu(model), "EPOCHS_RUN": int(next_epoch)}
    flat = getattr(model, "flat", None)
    if flat is not None and hasattr(flat, "exp_avg"):
        snap["OPTIM_STATE"] = {"step": flat.step, "exp_avg": flat.exp_avg.cpu(), "exp_avg_sq": flat.exp_avg_sq.cpu(),
                               "names": list(flat.names), "offsets": dict(flat.offsets)}
    snap["RNG_STATE"] = torch.get_rng_state()
    if extra:
        snap.update(extra)
    _atomic_save(snap, path)
    return True


def load_snapshot(path: str, model: torch.nn.Module, map_location="cpu", strict: bool = True) -> Dict[str, Any]:
    snap = torch.load(path, map_location=map_location, weights_only=False)
    state = snap["MODEL_STATE"] if isinstance(snap, dict) and "MODEL_STATE" in snap else snap
    model.load_state_dict(state, strict=strict)
    flat = getattr(model, "flat", None)
    if flat is not None:
        flat.rebind()
        opt = snap.get("OPTIM_STATE") if isinstance(snap, dict) else None
        if opt is not None and list(opt.get("names", [])) == list(flat.names):
            flat.step = int(opt["step"])
            flat.exp_avg.copy_(opt["exp_avg"])
            flat.exp_avg_sq.copy_(opt["exp_avg_sq"])
    if hasattr(model, "on_weights_changed"):
        model.on_weights_changed()
    fused = getattr(model, "_fused", None)
    if fused is not None and hasattr(fused, "sync_step"):
        fused.sync_step()                    # a pipelined optimiser keeps the Adam step count on the device
    out = {"EPOCHS_RUN": int(snap.get("EPOCHS_RUN", 0)) if isinstance(snap, dict) else 0}
    for k in ("ROUND", "RNG_STATE"):
        if isinstance(snap, dict) and k in snap:
            out[k] = snap[k]
    return out


def save_state_dict(path: str, model: torch.nn.Module) -> None:
    _atomic_save(model_state_fp32_cpu(model), path)


def load_state_dict_file(path: str) -> Dict[str, torch.Tensor]:
    return torch.load(path, map_location="cpu", weights_only=True)
