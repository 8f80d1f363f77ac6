"""Counter-based uniform RNG shared by the CPU oracle and the CUDA kernels.

Philox4x32-10 keyed by ``(seed, step)`` with counter ``(element_index // 4, stream, 0, 0)``; lane
``element_index % 4`` of the 128-bit output is the element's draw, and ``u = (x >> 8) * 2**-24``
lies in [0, 1).  Because the draw depends only on (seed, step, stream, global element index) the
quantisation of an element does not depend on launch geometry, bucket order or rank count, which
is what makes the fused kernels bit-comparable with the oracle.  The
reference draws ``torch.empty_like(t).uniform_()`` (``qsgd.py:23``), which cannot be reproduced
across devices.
"""
from __future__ import annotations

import numpy as np
import torch

PHILOX_M0 = 0xD2511F53
PHILOX_M1 = 0xCD9E8D57
PHILOX_W0 = 0x9E3779B9
PHILOX_W1 = 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32(c0, c1, c2, c3, k0, k1, rounds: int = 10):
    """Vectorised Philox4x32 over numpy uint64 arrays holding 32-bit values."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) & _MASK for c in (c0, c1, c2, c3))
    k0 = np.uint64(k0 & 0xFFFFFFFF)
    k1 = np.uint64(k1 & 0xFFFFFFFF)
    m0, m1 = np.uint64(PHILOX_M0), np.uint64(PHILOX_M1)
    for _ in range(rounds):
        p0 = m0 * c0
        p1 = m1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & _MASK, lo1, (hi0 ^ c3 ^ k1) & _MASK, lo0
        k0 = (k0 + np.uint64(PHILOX_W0)) & _MASK
        k1 = (k1 + np.uint64(PHILOX_W1)) & _MASK
    return c0, c1, c2, c3


def _uniform_of_counters(ctr: np.ndarray, seed: int, step: int, stream: int) -> np.ndarray:
    """The four 24-bit uniforms of each 64-bit counter: float32 array of shape (len(ctr), 4)."""
    zeros = np.zeros_like(ctr)
    r = philox4x32(ctr & _MASK, zeros + np.uint64(stream & 0xFFFFFFFF), ctr >> np.uint64(32), zeros,
                   seed & 0xFFFFFFFF, step & 0xFFFFFFFF)
    x = np.stack(r, axis=-1)
    return (x >> np.uint64(8)).astype(np.float32) * np.float32(2.0 ** -24)


def uniform_for_indices(index: torch.Tensor, seed: int, step: int, stream: int = 0) -> torch.Tensor:
    """u in [0,1) (float32, 24-bit) for each global element index (any int tensor).  Element i uses lane i & 3 of the
    Philox block with counter i >> 2, so neighbouring indices share a block: each distinct counter is evaluated once."""
    idx = index.detach().cpu().numpy().astype(np.uint64)
    shape = idx.shape
    idx = idx.reshape(-1)
    if idx.size == 0:
        return torch.zeros(shape, dtype=torch.float32)
    ctr = idx >> np.uint64(2)
    lane = (idx & np.uint64(3)).astype(np.int64)
    if idx.size > 1 and bool(np.all(ctr[1:] >= ctr[:-1])):          # sorted indices (top-k lists, ranges): one linear pass
        first = np.empty(ctr.size, dtype=bool)
        first[0] = True
        np.not_equal(ctr[1:], ctr[:-1], out=first[1:])
        uniq = ctr[first]
        inverse = np.cumsum(first) - 1
    else:
        uniq, inverse = np.unique(ctr, return_inverse=True)
    u4 = _uniform_of_counters(uniq, seed, step, stream)
    u = u4[inverse, lane]
    return torch.from_numpy(np.ascontiguousarray(u)).view(shape)


def uniform_like(t: torch.Tensor, seed: int, step: int, stream: int = 0, offset: int = 0) -> torch.Tensor:
    """Uniforms for a tensor whose first element has global index ``offset``: the blocks of the counters
    offset >> 2 .. (offset + n - 1) >> 2, laid end to end and cut to the range."""
    n = t.numel()
    if n == 0:
        return torch.zeros(t.shape, dtype=torch.float32)
    c0, c1 = offset >> 2, (offset + n - 1) >> 2
    u4 = _uniform_of_counters(np.arange(c0, c1 + 1, dtype=np.uint64), seed, step, stream)
    lead = offset & 3
    return torch.from_numpy(np.ascontiguousarray(u4.reshape(-1)[lead:lead + n])).view(t.shape)
