"""Top-k sparsification followed by QSGD (the report's methods 5 and 6).

Order of operations as described at ``README.md:20`` / ``Report.zip!main.tex:117`` and coded (but
commented out) at ``src/Compresssor/qsgd.py:14-15``: sparsify, densify, then quantise the sparse
tensor -- so the L2 norm is that of the *k kept values*.  The payload keeps only the k entries.
"""
from __future__ import annotations

from typing import Optional

import torch

from .base import Compressor
from .qsgd import code_width, level_dtype, qsgd_dequantize, qsgd_quantize
from .rng import uniform_for_indices
from .topk import sparsify


class TopKQSGDCompressor(Compressor):
    name = "topk-qsgd"

    def __init__(self, compress_ratio: float = 0.5, quantum_num: int = 128, seed: int = 0):
        self.compress_ratio = float(compress_ratio)
        self.quantum_num = int(quantum_num)
        self.seed = int(seed)
        self.step = 0

    def compress(self, tensor: torch.Tensor, uniform: Optional[torch.Tensor] = None, offset: int = 0):
        values, indices = sparsify(tensor, self.compress_ratio)
        if uniform is not None:
            u = uniform.reshape(-1)[indices]
        else:
            u = uniform_for_indices(indices + offset, self.seed, self.step).to(values.device)
        q, norm = qsgd_quantize(values, self.quantum_num, u)       # norm over the kept values
        q = q.to(level_dtype(self.quantum_num))
        return (q, indices.to(torch.int32)), (norm, tensor.numel(), tensor.size())

    def decompress(self, payload, ctx):
        q, indices = payload
        norm, numel, shape = ctx
        vals = qsgd_dequantize(q, norm, self.quantum_num)
        out = torch.zeros(numel, dtype=torch.float32, device=vals.device)
        out[indices.long()] = vals
        return out.view(shape)

    def payload_nbytes(self, payload, ctx) -> int:
        q, _ = payload
        numel = ctx[1]
        nnz = int((q != 0).sum())
        cw = code_width(self.quantum_num, sparse=True)
        idx_bytes = 2 * nnz if q.numel() * 16 <= numel else (numel + 7) // 8
        return idx_bytes + (nnz * cw + 7) // 8 + 4
