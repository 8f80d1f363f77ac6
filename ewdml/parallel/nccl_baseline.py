"""``--backend nccl``: the same algorithms on stock PyTorch ops + NCCL collectives (the measured baseline).

BASELINE.json: "a path that only calls NCCL or MPI for the gradient exchange is the baseline, not the
product".  These classes are that path behind the same ``Exchange`` protocol as the gloo plumbing
(:mod:`ewdml.parallel.host_exchange`) and the native kernels (:mod:`ewdml.parallel.native_exchange`):
the reference's compressor math as torch ops per parameter tensor on the GPU (``torch.topk`` on |g|,
``Compresssor/TopK.py:5-10``; QSGD levels with ``torch.rand``, ``Compresssor/qsgd.py:17-28``), fixed-size
(levels int16, indices int32, norms fp32) buffers moved by ONE ``all_gather`` / ``gather`` per buffer per
step, decompression with ``index_add_``, then the external-gradient optimizer.  None of this repo's CUDA
kernels run here.  ``bench.py --impl nccl`` (``baseline/nccl_same_algorithm.py``) times the same thing with
a stock ``torch.optim.SGD``.
"""
from __future__ import annotations

import time
from typing import List

import torch
import torch.distributed as dist

from .flat import FlatModel
from .host_exchange import Exchange, ExchangeStats


class _TorchCodec:
    """Per-tensor top-k -> QSGD into fixed-size device buffers; ``compressor`` in none|qsgd|topk|topk-qsgd."""

    def __init__(self, flat: FlatModel, compressor: str, levels: int, ratio: float):
        self.flat, self.comp, self.s = flat, compressor, float(levels)
        self.topk = compressor in ("topk", "topk-qsgd")
        self.quant = compressor in ("qsgd", "topk-qsgd")
        self.ks = [max(1, int(s.numel * ratio)) if self.topk else s.numel for s in flat.segments]
        self.offs = [0]
        for k in self.ks:
            self.offs.append(self.offs[-1] + k)
        self.K, self.nseg = self.offs[-1], len(flat.segments)
        dev = flat.param.device
        self.val = torch.empty(self.K, dtype=torch.int16 if self.quant else torch.float32, device=dev)
        self.idx = torch.empty(self.K if self.topk else 0, dtype=torch.int32, device=dev)
        self.norm = torch.zeros(self.nseg, dtype=torch.float32, device=dev)

    @property
    def nbytes(self) -> int:
        return self.val.numel() * self.val.element_size() + self.idx.numel() * 4 + self.norm.numel() * 4

    def encode(self, grad: torch.Tensor) -> None:
        for i, seg in enumerate(self.flat.segments):
            g = grad[seg.offset:seg.offset + seg.numel]
            lo, hi = self.offs[i], self.offs[i + 1]
            if self.topk:
                _, ix = torch.topk(g.abs(), self.ks[i], sorted=False)
                v = g[ix]
                self.idx[lo:hi] = ix.to(torch.int32)
            else:
                v = g
            if self.quant:
                norm = v.norm()
                lf = torch.where(norm > 0, self.s / norm, torch.zeros_like(norm)) * v.abs()   # zero-norm tensors encode to zeros
                prev = lf.floor()
                level = (prev + (torch.rand_like(v) < (lf - prev))).clamp_(max=self.s)
                self.val[lo:hi] = (level * v.sign()).to(torch.int16)
                self.norm[i] = norm
            else:
                self.val[lo:hi] = v

    def decode_into(self, out: torch.Tensor, val: torch.Tensor, idx: torch.Tensor, norm: torch.Tensor) -> None:
        """out (flat fp32) += decoded payload."""
        for i, seg in enumerate(self.flat.segments):
            lo, hi = self.offs[i], self.offs[i + 1]
            v = val[lo:hi].float()
            if self.quant:
                v = v * (norm[i] / self.s)
            dst = out[seg.offset:seg.offset + seg.numel]
            if self.topk:
                dst.index_add_(0, idx[lo:hi].long(), v)
            else:
                dst.add_(v)


def _u8(t: torch.Tensor) -> torch.Tensor:
    return t.view(torch.uint8)          # NCCL has no int16


def _all_gather_flat(out: torch.Tensor, inp: torch.Tensor, group) -> None:
    if inp.is_cuda:
        dist.all_gather_into_tensor(out, inp, group=group)
    else:
        dist.all_gather(list(out.chunk(dist.get_world_size(group))), inp, group=group)


class _NcclBase(Exchange):
    def __init__(self, flat, cfg, optimizer, group=None):
        super().__init__(flat, cfg, optimizer, group)
        # (the classes are device-agnostic torch code; on CPU tensors over gloo they serve as a plumbing test)
        self.codec = _TorchCodec(flat, cfg.resolved_compressor(), cfg.qsgd_levels, cfg.topk_ratio)
        self.avg = torch.zeros(flat.total, dtype=torch.float32, device=flat.param.device)

    def _apply(self, dense: torch.Tensor) -> None:
        self.optimizer.step(grads=[self.flat.view(dense, i) for i in range(len(self.flat.segments))])


class NcclAllgather(_NcclBase):
    def __init__(self, flat, cfg, optimizer, group=None):
        super().__init__(flat, cfg, optimizer, group)
        c, W = self.codec, self.world
        self.g_val = torch.empty(W * c.val.numel(), dtype=c.val.dtype, device=c.val.device)
        self.g_idx = torch.empty(W * c.idx.numel(), dtype=torch.int32, device=c.val.device)
        self.g_norm = torch.empty(W * c.nseg, dtype=torch.float32, device=c.val.device)

    def step(self, step_idx: int, is_sync_step: bool = True, want_stats: bool = False) -> ExchangeStats:
        if not is_sync_step:
            return self.local_step()
        st, c, W = ExchangeStats(), self.codec, self.world
        t0 = time.perf_counter()
        c.encode(self.flat.grad)
        if W > 1:
            _all_gather_flat(_u8(self.g_val), _u8(c.val), self.group)
            if c.idx.numel():
                _all_gather_flat(self.g_idx, c.idx, self.group)
            _all_gather_flat(self.g_norm, c.norm, self.group)
        else:
            self.g_val.copy_(c.val); self.g_idx.copy_(c.idx); self.g_norm.copy_(c.norm)
        st.t_send = time.perf_counter() - t0
        t0 = time.perf_counter()
        self.avg.zero_()
        nv, ni = c.val.numel(), c.idx.numel()
        for r in range(W):
            c.decode_into(self.avg, self.g_val[r * nv:(r + 1) * nv], self.g_idx[r * ni:(r + 1) * ni],
                          self.g_norm[r * c.nseg:(r + 1) * c.nseg])
        self.avg.div_(W)
        self._apply(self.avg)
        st.t_recv = time.perf_counter() - t0
        st.bytes_sent = st.bytes_recv = (W - 1) * c.nbytes
        return st


class NcclAllreduce(_NcclBase):
    def step(self, step_idx: int, is_sync_step: bool = True, want_stats: bool = False) -> ExchangeStats:
        if not is_sync_step:
            return self.local_step()
        st = ExchangeStats()
        t0 = time.perf_counter()
        if self.world > 1:
            dist.all_reduce(self.flat.grad, group=self.group)
            self.flat.grad.div_(self.world)
        st.t_send = time.perf_counter() - t0
        self.optimizer.step(grads=self.flat.grad_views())
        st.bytes_sent = st.bytes_recv = int(2 * (self.world - 1) / max(1, self.world) * self.flat.dense_nbytes())
        return st


class NcclPS(_NcclBase):
    """Rank 0 = server.  Push: ``gather`` of the three payload buffers; pull: ``broadcast`` of the weights
    (methods 1/2) or of the dense averaged gradient (methods 3-5; a library path gains nothing from
    re-compressing the pull, so it is sent dense)."""

    def __init__(self, flat, cfg, optimizer, group=None, role: str = "worker"):
        super().__init__(flat, cfg, optimizer, group)
        self.role, self.pull_mode = role, cfg.resolved_pull()
        c, W = self.codec, self.world
        dev = c.val.device
        mk = lambda t: [torch.empty_like(t) for _ in range(W)] if role == "server" else None
        self.l_val, self.l_idx, self.l_norm = mk(_u8(c.val)), mk(c.idx), mk(c.norm)
        self.dense = torch.zeros(flat.total, dtype=torch.float32, device=dev)

    def step(self, step_idx: int, is_sync_step: bool = True, want_stats: bool = False) -> ExchangeStats:
        if not is_sync_step:
            return self.local_step() if self.role == "worker" else ExchangeStats()
        st, c, W = ExchangeStats(), self.codec, self.world
        t0 = time.perf_counter()
        if self.role == "worker":
            c.encode(self.flat.grad)
        dist.gather(_u8(c.val), self.l_val, dst=0, group=self.group)
        if c.idx.numel():
            dist.gather(c.idx, self.l_idx, dst=0, group=self.group)
        dist.gather(c.norm, self.l_norm, dst=0, group=self.group)
        st.t_send = time.perf_counter() - t0
        t0 = time.perf_counter()
        if self.role == "server":
            self.avg.zero_()
            for r in range(1, W):
                c.decode_into(self.avg, self.l_val[r].view(c.val.dtype), self.l_idx[r] if c.idx.numel() else c.idx, self.l_norm[r])
            self.avg.div_(max(1, W - 1))
            if self.pull_mode == "weights":
                self._apply(self.avg)
        if self.pull_mode == "weights":
            dist.broadcast(self.flat.param, src=0, group=self.group)
            if self.flat.param_bf16 is not None:
                self.flat.param_bf16.copy_(self.flat.param)
        else:
            buf = self.avg if self.role == "server" else self.dense
            dist.broadcast(buf, src=0, group=self.group)
            if self.role == "worker":
                self._apply(buf)
        st.t_recv = time.perf_counter() - t0
        n_dense = self.flat.dense_nbytes()
        if self.role == "worker":
            st.bytes_sent, st.bytes_recv = c.nbytes, n_dense
        else:
            st.bytes_recv, st.bytes_sent = (W - 1) * c.nbytes, (W - 1) * n_dense
        return st


def build_nccl_exchange(role: str, flat: FlatModel, cfg, optimizer, group=None) -> Exchange:
    if cfg.exchange == "ps":
        return NcclPS(flat, cfg, optimizer, group, role=role)
    if cfg.exchange == "allgather":
        return NcclAllgather(flat, cfg, optimizer, group)
    return NcclAllreduce(flat, cfg, optimizer, group)
