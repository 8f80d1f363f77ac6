"""Top-k magnitude sparsification -- CPU/torch oracle and user-facing compressor.

Semantics of ``src/Compresssor/TopK.py:5-35``: flatten, ``k = max(1, int(numel*ratio))``, keep the k
largest |values| with their signs and flat indices; decompress scatters into zeros and restores
the shape.  Ties at the threshold: the oracle keeps the lowest indices (deterministic); the CUDA
kernel keeps an arbitrary subset of the tied elements of the same size, so dense results agree
whenever the k-th magnitude is unique.
"""
from __future__ import annotations

from typing import Tuple

import torch

from .base import Compressor


def topk_count(numel: int, ratio: float) -> int:
    return max(1, int(numel * ratio))


def sparsify(tensor: torch.Tensor, compress_ratio: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """(values, flat indices) of the k largest magnitudes; the indices come back in ASCENDING order (the host wire format
    stores them that way and relies on it)."""
    flat = tensor.detach().reshape(-1)
    k = topk_count(flat.numel(), compress_ratio)
    mag = flat.abs()
    n = mag.numel()
    if k >= n:
        order = torch.arange(n, device=flat.device)
        return flat[order], order
    # Selection instead of a full sort: the k-th largest magnitude is the threshold, everything above it is kept and the
    # remaining slots go to the LOWEST indices among the elements equal to it -- the same set a stable descending sort
    # would give, in O(n).
    if bool(torch.isnan(mag).any()):
        # NaNs order above every number and compare false with everything: keep the sort's treatment of them
        order = torch.sort(mag, descending=True, stable=True).indices[:k]
        order = torch.sort(order).values
        return flat[order], order
    thr = torch.kthvalue(mag, n - k + 1).values
    keep = mag > thr
    need = k - int(keep.sum())
    if need > 0:
        ties = torch.nonzero(mag == thr).flatten()[:need]
        keep[ties] = True
    order = torch.nonzero(keep).flatten()
    return flat[order], order


def desparsify(tensors, numel: int) -> torch.Tensor:
    values, indices = tensors
    out = torch.zeros(numel, dtype=values.dtype, device=values.device)
    out[indices.long()] = values
    return out


class TopKCompressor(Compressor):
    name = "topk"

    def __init__(self, compress_ratio: float = 0.5):
        self.compress_ratio = float(compress_ratio)

    def compress(self, tensor: torch.Tensor):
        values, indices = sparsify(tensor, self.compress_ratio)
        return (values, indices.to(torch.int32)), (tensor.numel(), tensor.size())

    def decompress(self, tensors, ctx):
        numel, shape = ctx
        return desparsify(tensors, numel).view(shape)

    def payload_nbytes(self, tensors, ctx=None) -> int:
        values, _ = tensors
        k = values.numel()
        numel = ctx[0] if ctx is not None else k
        # index list (u16 tile-local) vs bitmap, whichever the wire format would pick
        idx_bytes = 2 * k if k * 16 <= numel else (numel + 7) // 8
        return idx_bytes + 4 * k
