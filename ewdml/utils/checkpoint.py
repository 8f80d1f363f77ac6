"""Checkpoint save / load / resume.

Format parity: the model file is a plain ``torch.save(state_dict)`` zip container
with the reference's key names, so ``torch.load`` works both ways -- the reference evaluator
(``distributed_evaluator.py:104-107``) can read our files and we can read ``src/model_step_.zip``.
File naming follows ``distributed_worker.py:392-398`` (``train_dir + "model_step_"``, same file
overwritten) when ``legacy_name`` is set, otherwise the master's ``model_step_<step>`` variant
(``sync_replicas_master_nn.py:243-244``).

Fixes: one writer (callers pass ``is_writer``), atomic tmp+rename instead of every worker racing on
the same path, and an optional sidecar (``<path>.state``) holding optimizer state, step counter,
RNG state and the compressor's error-feedback residual so training can ``--resume``.
"""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Any, Dict, Optional

import torch


def model_path(train_dir: str, step: Optional[int] = None, legacy_name: bool = True) -> str:
    base = train_dir + "model_step_"        # string concatenation, exactly like the reference
    return base if (legacy_name or step is None) else base + str(step)


def _atomic_save(obj: Any, path: str) -> None:
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    tmp = f"{path}.tmp.{os.getpid()}"
    with open(tmp, "wb") as f:
        torch.save(obj, f)
    os.replace(tmp, path)


def cpu_state_dict(model: torch.nn.Module) -> "OrderedDict[str, torch.Tensor]":
    """Contiguous CPU copy of the state_dict with metadata preserved."""
    sd = model.state_dict()
    out = OrderedDict((k, v.detach().to("cpu").contiguous().clone()) for k, v in sd.items())
    if hasattr(sd, "_metadata"):
        out._metadata = sd._metadata  # type: ignore[attr-defined]
    return out


def save_model(model: torch.nn.Module, path: str, is_writer: bool = True) -> None:
    if is_writer:
        _atomic_save(cpu_state_dict(model), path)


def save_sidecar(path: str, *, step: int, optimizer_state: Optional[Dict] = None,
                 extra: Optional[Dict] = None, is_writer: bool = True) -> None:
    if not is_writer:
        return
    blob = {"format": "ewdml-sidecar-1", "step": int(step), "optimizer": optimizer_state,
            "torch_rng": torch.get_rng_state(), "extra": extra or {}}
    _atomic_save(blob, path + ".state")


def load_model(model: torch.nn.Module, path: str, strict: bool = True) -> None:
    with open(path, "rb") as f:
        sd = torch.load(f, map_location="cpu", weights_only=True)
    model.load_state_dict(sd, strict=strict)


def load_sidecar(path: str) -> Optional[Dict]:
    p = path + ".state"
    if not os.path.isfile(p):
        return None
    with open(p, "rb") as f:
        return torch.load(f, map_location="cpu", weights_only=False)
