"""QSGD stochastic quantisation -- CPU/torch oracle and user-facing compressor.

Math as in ``src/Compresssor/qsgd.py:12-40`` and ``horovod_compression.py:16-43``:
``norm = ||t||_2`` over the whole tensor, ``l = s*|t|/norm``, ``level = floor(l) + [u < l-floor(l)]``
with ``u ~ U[0,1)``, payload ``(level*sign(t), norm)``, decode ``norm/s * q``.

Differences: a zero-norm tensor quantises to all-zero levels instead of
dividing by zero; levels are clamped to ``s``; the uniforms come from a counter-based generator
(:mod:`ewdml.compress.rng`) or are injected, so results are reproducible and comparable with the
CUDA kernels; the payload dtype is the smallest signed integer that holds ``[-s, s]`` instead of
float32 (the reference never packs, ``qsgd.py:27-32``).
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch

from .base import Compressor
from .rng import uniform_like


def level_dtype(s: int) -> torch.dtype:
    return torch.int8 if s <= 127 else (torch.int16 if s <= 32767 else torch.int32)


def code_width(s: int, sparse: bool) -> int:
    """Bits per transmitted element in the packed wire format.

    Sparse payloads (top-k list or bitmap) never carry level 0, so they store ``level-1`` in
    ``ceil(log2 s)`` bits plus a sign bit; s=128 -> 8 bits.
    """
    mag = max(1, math.ceil(math.log2(s))) if sparse else max(1, math.ceil(math.log2(s + 1)))
    return mag + 1


def qsgd_quantize(t: torch.Tensor, s: int, u: torch.Tensor, norm: Optional[torch.Tensor] = None
                  ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Signed integer levels (float tensor holding integers) and the norm used."""
    t = t.detach().to(torch.float32)
    if norm is None:
        norm = t.norm()
    norm = norm.reshape(()).to(torch.float32)
    if t.is_cuda:
        # device tensors: no host branch (float(norm) would synchronise once per tensor) -- a zero / non-finite norm selects
        # scale 0, every level becomes 0 and the norm that travels is 0, on the tensor's own device
        ok = torch.isfinite(norm) & (norm > 0)
        scale = torch.where(ok, float(s) / torch.where(ok, norm, torch.ones_like(norm)), torch.zeros_like(norm))
        lf = scale * t.abs()
        prev = lf.floor()
        level = (prev + (u.to(torch.float32) < (lf - prev)).to(torch.float32)).clamp_(max=float(s))
        return level * t.sign(), torch.where(ok, norm, torch.zeros_like(norm)).reshape(1)
    if float(norm) == 0.0 or not math.isfinite(float(norm)):
        # zero tensor, or a sum of squares that overflows fp32: nothing is sent and the receiver reconstructs zeros
        # (the reference divides by the norm unguarded, qsgd.py:21; a non-finite norm would decode to NaN)
        return torch.zeros_like(t), torch.zeros(1, dtype=torch.float32, device=t.device)
    scale = torch.tensor(float(s), dtype=torch.float32) / norm      # s / norm first, as the reference
    lf = scale * t.abs()
    prev = lf.floor()
    level = prev + (u.to(torch.float32) < (lf - prev)).to(torch.float32)
    level = level.clamp_(max=float(s))
    return level * t.sign(), norm.reshape(1)


def qsgd_dequantize(q: torch.Tensor, norm: torch.Tensor, s: int) -> torch.Tensor:
    return (norm.reshape(()).to(torch.float32) / float(s)) * q.to(torch.float32)


class QSGDCompressor(Compressor):
    """``compress(t) -> (q, norm)``; ``decompress((q, norm))`` or ``decompress(q, norm)``.

    ``topk_ratio`` enables the report's method 5 (top-k sparsification *before* quantisation,
    ``qsgd.py:14-15``); see :class:`ewdml.compress.topk_qsgd.TopKQSGDCompressor` for the packed form.
    """

    name = "qsgd"

    def __init__(self, quantum_num: int = 128, seed: int = 0, pack: bool = True):
        self.quantum_num = int(quantum_num)
        self.seed = int(seed)
        self.step = 0
        self.pack = pack

    # -- uniforms ---------------------------------------------------------------------
    def _uniform(self, t: torch.Tensor, u: Optional[torch.Tensor], offset: int) -> torch.Tensor:
        if u is not None:
            return u.reshape(t.shape)
        return uniform_like(t, self.seed, self.step, stream=0, offset=offset).to(t.device)

    def compress(self, tensor: torch.Tensor, uniform: Optional[torch.Tensor] = None, offset: int = 0):
        q, norm = qsgd_quantize(tensor, self.quantum_num, self._uniform(tensor, uniform, offset))
        if self.pack:
            q = q.to(level_dtype(self.quantum_num))
        return q, norm

    def decompress(self, payload, ctx=None):
        if ctx is None:
            payload, ctx = payload          # parameter-server flavour: one tuple argument
        return qsgd_dequantize(payload, ctx, self.quantum_num)

    def payload_nbytes(self, payload, ctx=None) -> int:
        if ctx is None:
            payload, ctx = payload
        nnz = int((payload != 0).sum())
        n = payload.numel()
        cw = code_width(self.quantum_num, sparse=True)
        # bitmap of non-zero levels + packed codes + norm (the CUDA "BITMAP" wire mode)
        return (n + 7) // 8 + (nnz * cw + 7) // 8 + 4


class HorovodQSGDCompressor(Compressor):
    """Static-method flavour of ``horovod_compression.py:11-43`` (s fixed at 128)."""

    name = "qsgd-hvd"
    quantum_num = 128
    _step = 0

    @staticmethod
    def compress(tensor):
        u = torch.rand(tensor.shape, dtype=torch.float32, device=tensor.device)
        q, norm = qsgd_quantize(tensor, HorovodQSGDCompressor.quantum_num, u)
        return q, norm

    @staticmethod
    def decompress(tensor_compressed, ctx):
        return qsgd_dequantize(tensor_compressed, ctx, HorovodQSGDCompressor.quantum_num)

    def payload_nbytes(self, payload, ctx=None) -> int:
        return payload.numel() + 4
