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


def _file_crc(path: str) -> int:
    import zlib
    crc = 0
    with open(path, "rb") as f:
        while True:
            chunk = f.read(1 << 20)
            if not chunk:
                break
            crc = zlib.crc32(chunk, crc)
    return crc & 0xFFFFFFFF


def _atomic_save(obj: Any, path: str) -> int:
    """tmp + rename; returns the CRC-32 of the bytes written."""
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    tmp = f"{path}.tmp.{os.getpid()}"
    with open(tmp, "wb") as f:
        torch.save(obj, f)
    crc = _file_crc(tmp)
    os.replace(tmp, path)
    return crc


def cpu_state_dict(model: torch.nn.Module) -> "OrderedDict[str, torch.Tensor]":
    """Contiguous CPU copy of the state_dict with metadata preserved."""
    sd = model.state_dict()
    out = OrderedDict((k, v.detach().to("cpu").contiguous().clone()) for k, v in sd.items())
    if hasattr(sd, "_metadata"):
        out._metadata = sd._metadata  # type: ignore[attr-defined]
    return out


def save_model(model: torch.nn.Module, path: str, is_writer: bool = True) -> Optional[int]:
    """Writes the reference-format model file; returns its CRC-32 (pass it to :func:`save_sidecar`)."""
    if is_writer:
        return _atomic_save(cpu_state_dict(model), path)
    return None


def save_sidecar(path: str, *, step: int, optimizer_state: Optional[Dict] = None,
                 extra: Optional[Dict] = None, is_writer: bool = True, model_crc: Optional[int] = None,
                 rng_rank: Optional[int] = None) -> None:
    """Model file first, sidecar second.  The two are separate renames, so a crash in between leaves a NEW model beside an
    OLD sidecar; ``model_crc`` ties the sidecar to the exact model bytes it was written for and :func:`load_sidecar` ignores
    a sidecar whose model has changed (weights are kept, step / optimizer state restart).  ``rng_rank``: whose generator
    state ``torch_rng`` is -- other ranks must not adopt it (they would all draw the same dropout / QSGD streams)."""
    if not is_writer:
        return
    blob = {"format": "ewdml-sidecar-1", "step": int(step), "optimizer": optimizer_state,
            "torch_rng": torch.get_rng_state(), "rng_rank": rng_rank, "model_crc": model_crc, "extra": extra or {}}
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
        side = torch.load(f, map_location="cpu", weights_only=False)
    crc = side.get("model_crc") if isinstance(side, dict) else None
    if crc is not None and os.path.isfile(path) and _file_crc(path) != crc:
        import logging
        logging.getLogger("ewdml").warning("%s was written for a different version of %s (interrupted checkpoint?): "
                                           "keeping the weights, ignoring the stale step / optimizer state", p, path)
        return None
    return side


def reseed_for_rank(seed: int, rank: int, step: int) -> None:
    """Generator state of a rank that did not write the sidecar: derived from (seed, rank, step) so that ranks keep
    distinct dropout / quantisation streams after a resume."""
    torch.manual_seed((int(seed) + 1000 * (int(rank) + 1)) * 1_000_003 + int(step))
