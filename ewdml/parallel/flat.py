"""Flat parameter / gradient / momentum storage with a per-tensor segment table.

The reference exchanges one collective per parameter tensor (``distributed_worker.py:317-351``,
``sync_replicas_master_nn.py:218-232``) and keeps per-tensor receive buffers
(``ModelBuffer``, ``distributed_worker.py:41-59``; ``GradientAccumulator``,
``sync_replicas_master_nn.py:49-68``).  Here every parameter is a *view* into one flat fp32
buffer (likewise its gradient and momentum), and the per-tensor semantics of the compressors
(norm and top-k per tensor) are carried by a segment table.  One kernel launch then covers all
tensors, and on the GPU the gradient buffer lives in the symmetric peer heap so collectives need
no staging copy.

Layout rules: segment starts are aligned to ``align`` elements (default 32 -> 128 B, so bitmap
words and 16-byte vector accesses never straddle tensors); 4-D conv weights are stored in
channels-last order ``[C_out][kh][kw][C_in]`` when ``channels_last`` is set -- the K-major operand
layout of the implicit-GEMM kernels -- while keeping their logical ``[C_out, C_in, kh, kw]`` shape
(state_dict compatible).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, List, Optional

import torch


@dataclass
class Segment:
    name: str
    offset: int          # element offset in the flat buffers
    numel: int
    shape: torch.Size
    channels_last: bool


def _round_up(x: int, a: int) -> int:
    return (x + a - 1) // a * a


class FlatModel:
    def __init__(self, model: torch.nn.Module, device: Optional[torch.device] = None,
                 channels_last: bool = False, align: int = 32, with_momentum: bool = True,
                 with_bf16: bool = False,
                 allocator: Optional[Callable[[str, int, torch.dtype], torch.Tensor]] = None):
        self.model = model
        self.device = device or next(model.parameters()).device
        self.align = align
        params = [(n, p) for n, p in model.named_parameters()]
        self.params = [p for _, p in params]
        self.segments: List[Segment] = []
        off = 0
        for n, p in params:
            cl = bool(channels_last and p.dim() == 4)
            self.segments.append(Segment(n, off, p.numel(), p.shape, cl))
            off = _round_up(off + p.numel(), align)
        self.total = max(off, align)
        alloc = allocator or (lambda tag, n, dt: torch.zeros(n, dtype=dt, device=self.device))
        self.param = alloc("param", self.total, torch.float32)
        self.grad = alloc("grad", self.total, torch.float32)
        self.momentum = alloc("momentum", self.total, torch.float32) if with_momentum else None
        self.param_bf16 = alloc("param_bf16", self.total, torch.bfloat16) if with_bf16 else None
        self.momentum_initialised = False
        with torch.no_grad():
            for seg, p in zip(self.segments, self.params):
                view = self._view(self.param, seg)
                view.copy_(p.detach().to(self.device))
                p.data = view
                p.grad = self._view(self.grad, seg)
        self.grad.zero_()
        # move the buffers (BN running stats) too
        for b in model.buffers():
            b.data = b.data.to(self.device)
        if self.param_bf16 is not None:
            self.param_bf16.copy_(self.param)
            for seg, p in zip(self.segments, self.params):
                # storage-order bf16 copy of this parameter, kept current by the fused optimizer kernel
                p._ewdml_bf16 = self.param_bf16[seg.offset:seg.offset + seg.numel]

    # ------------------------------------------------------------------
    def _view(self, flat: torch.Tensor, seg: Segment) -> torch.Tensor:
        v = flat[seg.offset:seg.offset + seg.numel]
        if seg.channels_last:
            co, ci, kh, kw = seg.shape
            return v.view(co, kh, kw, ci).permute(0, 3, 1, 2)
        return v.view(seg.shape)

    def view(self, flat: torch.Tensor, i: int) -> torch.Tensor:
        return self._view(flat, self.segments[i])

    def grad_views(self) -> List[torch.Tensor]:
        return [self._view(self.grad, s) for s in self.segments]

    def flat_slices(self, flat: torch.Tensor) -> List[torch.Tensor]:
        """1-D storage-order slices per segment (what the kernels and compressors see)."""
        return [flat[s.offset:s.offset + s.numel] for s in self.segments]

    def zero_grad(self) -> None:
        self.grad.zero_()
        for seg, p in zip(self.segments, self.params):
            if p.grad is None or p.grad.data_ptr() != self.grad.data_ptr() + 4 * seg.offset:
                p.grad = self._view(self.grad, seg)

    @property
    def num_params(self) -> int:
        return sum(s.numel for s in self.segments)

    def dense_nbytes(self) -> int:
        return 4 * self.num_params

    # BN running stats etc. are never exchanged (distributed_worker.py:294)
    def load_flat_params(self, flat: torch.Tensor) -> None:
        self.param.copy_(flat)
        if self.param_bf16 is not None:
            self.param_bf16.copy_(self.param)
