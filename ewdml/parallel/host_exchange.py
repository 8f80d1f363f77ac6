"""Gradient exchange strategies over ``torch.distributed`` host collectives (gloo / NCCL baseline).

This is the *plumbing* implementation of the three exchange patterns; the product path for B200
is :mod:`ewdml.parallel.native_exchange` (hand-written NVLink kernels).  Both expose the same
``Exchange`` protocol so the training loop does not care.

* :class:`PSWorker` / :class:`PSServer` -- synchronous parameter server, rank 0 = server
  (``sync_replicas_master_nn.py:158-232`` + ``distributed_worker.py:162-351``): workers *gather*
  compressed gradients to the server, the server decompresses, averages over the workers and
  either updates its model and broadcasts **weights** (report method 1/2) or broadcasts the
  averaged **gradient**, optionally re-compressed (methods 3/4/5).  One collective per direction
  per step instead of one per tensor.
* :class:`AllgatherExchange` -- every rank is a worker; payloads are all-gathered and each rank
  decodes all of them with the sender's own norm (fixes the Horovod path's "average the levels,
  rescale by the local norm" inconsistency, ``horovod_compression.py:36-43``).
* :class:`AllreduceExchange` -- dense averaging allreduce (``hvd.allreduce(op=Average)``,
  ``horvod_pytorch.py:197-201``) with a gradient pre-divide factor.

Method 6 (exchange every ``sync_period``-th step, local SGD in between, ``distributed_worker.py:
197-209``) is handled by the ``is_sync_step`` argument of ``Exchange.step``.
"""
from __future__ import annotations

import time
from dataclasses import dataclass
from typing import List, Optional

import torch
import torch.distributed as dist

from .flat import FlatModel
from .host_codec import HostCodec


@dataclass
class ExchangeStats:
    bytes_sent: int = 0
    bytes_recv: int = 0
    t_send: float = 0.0
    t_recv: float = 0.0


def _codec(flat: FlatModel, cfg, rank: int, compressor: Optional[str] = None, segments=None) -> HostCodec:
    return HostCodec(segments if segments is not None else flat.segments, compressor or cfg.resolved_compressor(),
                     cfg.qsgd_levels, cfg.topk_ratio, seed=cfg.seed, rank=rank)


class Exchange:
    """Protocol: ``step(step_idx, is_sync_step)`` consumes ``flat.grad`` and updates ``flat.param``."""

    def __init__(self, flat: FlatModel, cfg, optimizer, group=None):
        self.flat, self.cfg, self.optimizer, self.group = flat, cfg, optimizer, group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        # error feedback (optional, off = reference semantics): what the compressor dropped is added to the next gradient
        ef = bool(getattr(cfg, "error_feedback", False)) and cfg.resolved_compressor() != "none"
        self.residual = torch.zeros(flat.total, dtype=torch.float32) if ef else None
        self._own = torch.zeros(flat.total, dtype=torch.float32) if ef else None

    def _ef_source(self) -> torch.Tensor:
        """The vector to compress: the gradient, plus the carried residual when error feedback is on."""
        if self.residual is None:
            return self.flat.grad
        self.residual.add_(self.flat.grad.detach().cpu())
        return self.residual

    def _ef_update(self, codec: HostCodec, msg: torch.Tensor) -> None:
        if self.residual is not None:
            codec.decode(msg, out=self._own, accumulate=False, total=self.flat.total)
            for seg in codec.segments:                      # only this codec's segments (bucketed exchanges)
                self.residual[seg.offset:seg.offset + seg.numel].sub_(self._own[seg.offset:seg.offset + seg.numel])

    def local_step(self) -> ExchangeStats:
        """Local SGD on this rank's own gradient (non-sync steps of method 6)."""
        self.optimizer.step(grads=self.flat.grad_views())
        return ExchangeStats()

    def step(self, step_idx: int, is_sync_step: bool = True) -> ExchangeStats:
        raise NotImplementedError

    def broadcast_initial_weights(self, src: int = 0) -> None:
        """Fix for the reference's missing initial weight sync (SURVEY.md 3.1 note)."""
        if self.world > 1:
            buf = self.flat.param.detach().cpu() if self.flat.param.is_cuda and _is_gloo(self.group) else self.flat.param
            dist.broadcast(buf, src=src, group=self.group)
            if buf is not self.flat.param:
                self.flat.param.copy_(buf)
            for b in self.flat.model.buffers():     # BN running stats start equal too
                t = b.detach().cpu() if b.is_cuda and _is_gloo(self.group) else b
                dist.broadcast(t, src=src, group=self.group)
                if t is not b:
                    b.copy_(t)


def _is_gloo(group) -> bool:
    return dist.get_backend(group) == "gloo"


# --------------------------------------------------------------------------------------
# Parameter server
# --------------------------------------------------------------------------------------
class PSWorker(Exchange):
    def __init__(self, flat, cfg, optimizer, group=None):
        super().__init__(flat, cfg, optimizer, group)
        self.push = _codec(flat, cfg, self.rank)
        self.pull_mode = cfg.resolved_pull()
        pull_comp = cfg.resolved_compressor() if cfg.resolved_compress_pull() else "none"
        self.pull = _codec(flat, cfg, 0, pull_comp)
        self._dense = torch.zeros(flat.total, dtype=torch.float32)
        self._wbuf = torch.zeros(flat.total, dtype=torch.float32)
        self._pull_msg = torch.empty(self.pull.nbytes, dtype=torch.uint8)

    def step(self, step_idx: int, is_sync_step: bool = True) -> ExchangeStats:
        if not is_sync_step:
            return self.local_step()
        st = ExchangeStats()
        t0 = time.perf_counter()
        msg = self.push.encode(self._ef_source(), step_idx)
        self._ef_update(self.push, msg)
        dist.gather(msg, None, dst=0, group=self.group)
        st.t_send = time.perf_counter() - t0
        st.bytes_sent = msg.numel()
        t0 = time.perf_counter()
        if self.pull_mode == "weights":
            dist.broadcast(self._wbuf, src=0, group=self.group)
            self.flat.load_flat_params(self._wbuf.to(self.flat.param.device))
            st.bytes_recv = self._wbuf.numel() * 4
        else:
            dist.broadcast(self._pull_msg, src=0, group=self.group)
            self.pull.decode(self._pull_msg, out=self._dense)
            st.bytes_recv = self._pull_msg.numel()
            dense = self._dense.to(self.flat.param.device)
            self.optimizer.step(grads=[self.flat.view(dense, i) for i in range(len(self.flat.segments))])
        st.t_recv = time.perf_counter() - t0
        return st


class PSServer(Exchange):
    """Rank 0.  Holds a model only when it has to (pull == weights), unlike the reference whose
    server builds a model it never uses in gradient-relay mode."""

    def __init__(self, flat, cfg, optimizer, group=None):
        super().__init__(flat, cfg, optimizer, group)
        self.num_workers = self.world - 1
        self.decoders = [_codec(flat, cfg, r) for r in range(self.world)]
        self.pull_mode = cfg.resolved_pull()
        pull_comp = cfg.resolved_compressor() if cfg.resolved_compress_pull() else "none"
        self.pull = _codec(flat, cfg, 0, pull_comp)
        n = self.decoders[0].nbytes
        self._slots = [torch.zeros(n, dtype=torch.uint8) for _ in range(self.world)]
        self._avg = torch.zeros(flat.total, dtype=torch.float32)

    def step(self, step_idx: int, is_sync_step: bool = True) -> ExchangeStats:
        if not is_sync_step:
            return ExchangeStats()      # workers run local SGD; nothing reaches the server
        st = ExchangeStats()
        t0 = time.perf_counter()
        dist.gather(self._slots[0], self._slots, dst=0, group=self.group)   # slot 0 = own dummy
        self._avg.zero_()
        for r in range(1, self.world):
            self.decoders[r].decode(self._slots[r], out=self._avg, accumulate=True)
        self._avg.div_(max(1, self.num_workers))
        st.t_recv = time.perf_counter() - t0
        st.bytes_recv = self.num_workers * self._slots[0].numel()
        t0 = time.perf_counter()
        if self.pull_mode == "weights":
            avg = self._avg.to(self.flat.param.device)
            self.optimizer.step(grads=[self.flat.view(avg, i) for i in range(len(self.flat.segments))])
            w = self.flat.param.detach().cpu()
            dist.broadcast(w, src=0, group=self.group)
            st.bytes_sent = self.num_workers * w.numel() * 4
        else:
            msg = self.pull.encode(self._avg, step_idx)
            dist.broadcast(msg, src=0, group=self.group)
            st.bytes_sent = self.num_workers * msg.numel()
        st.t_send = time.perf_counter() - t0
        return st


# --------------------------------------------------------------------------------------
# Server-less strategies
# --------------------------------------------------------------------------------------
class AllgatherExchange(Exchange):
    """``--bucket-size`` splits the model into contiguous buckets; with ``--overlap`` each bucket is encoded and its
    ``all_gather`` started (``async_op``) from the backward pass as soon as its gradients exist
    (:mod:`ewdml.parallel.overlap`), and ``finish_overlap`` waits, decodes and applies."""

    def __init__(self, flat, cfg, optimizer, group=None):
        super().__init__(flat, cfg, optimizer, group)
        from .overlap import partition_buckets
        self.buckets = partition_buckets(flat.segments, getattr(cfg, "bucket_size", 0))
        self.codecs = [[_codec(flat, cfg, r, segments=flat.segments[lo:hi]) for r in range(self.world)]
                       for lo, hi in self.buckets]
        self._slots = [[torch.zeros(c[0].nbytes, dtype=torch.uint8) for _ in range(self.world)] for c in self.codecs]
        self._avg = torch.zeros(flat.total, dtype=torch.float32)
        self._works, self._sent = [], 0

    def _launch(self, b: int, step_idx: int, async_op: bool, src: torch.Tensor):
        msg = self.codecs[b][self.rank].encode(src, step_idx)
        self._ef_update(self.codecs[b][self.rank], msg)
        self._sent += msg.numel()
        if self.world > 1:
            return dist.all_gather(self._slots[b], msg, group=self.group, async_op=async_op)
        self._slots[b][0].copy_(msg)
        return None

    def _apply_all(self) -> None:
        self._avg.zero_()
        for b in range(len(self.buckets)):
            for r in range(self.world):
                self.codecs[b][r].decode(self._slots[b][r], out=self._avg, accumulate=True, total=self.flat.total)
        self._avg.div_(self.world)
        avg = self._avg.to(self.flat.param.device)
        self.optimizer.step(grads=[self.flat.view(avg, i) for i in range(len(self.flat.segments))])

    def step(self, step_idx: int, is_sync_step: bool = True) -> ExchangeStats:
        if not is_sync_step:
            return self.local_step()
        st = ExchangeStats()
        t0 = time.perf_counter()
        self._sent = 0
        src = self._ef_source()
        for b in range(len(self.buckets) - 1, -1, -1):      # last layers first, the order backward produces them
            self._launch(b, step_idx, False, src)
        st.t_send = time.perf_counter() - t0
        t0 = time.perf_counter()
        self._apply_all()
        st.t_recv = time.perf_counter() - t0
        st.bytes_sent = st.bytes_recv = (self.world - 1) * self._sent
        return st

    # ---- backward-overlapped form ---------------------------------------------------------------------
    def begin_overlap(self, step_idx: int) -> None:
        if self.residual is not None:
            raise RuntimeError("--overlap and --error-feedback cannot be combined")
        self._works, self._sent, self._step = [], 0, step_idx

    def bucket_ready(self, b: int) -> None:
        self._works.append(self._launch(b, self._step, True, self.flat.grad))

    def finish_overlap(self) -> ExchangeStats:
        st = ExchangeStats()
        t0 = time.perf_counter()
        for w in self._works:
            if w is not None:
                w.wait()
        self._apply_all()
        st.t_recv = time.perf_counter() - t0
        st.bytes_sent = st.bytes_recv = (self.world - 1) * self._sent
        return st


class AllreduceExchange(Exchange):
    def __init__(self, flat, cfg, optimizer, group=None, predivide: float = 1.0):
        super().__init__(flat, cfg, optimizer, group)
        self.predivide = predivide

    def step(self, step_idx: int, is_sync_step: bool = True) -> ExchangeStats:
        if not is_sync_step:
            return self.local_step()
        st = ExchangeStats()
        t0 = time.perf_counter()
        g = self.flat.grad
        staged = g.detach().cpu() if (g.is_cuda and _is_gloo(self.group)) else g
        if self.predivide != 1.0:
            staged.div_(self.predivide)
        if self.world > 1:
            dist.all_reduce(staged, op=dist.ReduceOp.SUM, group=self.group)
        staged.mul_(self.predivide / self.world)
        if staged is not g:
            g.copy_(staged)
        st.t_send = time.perf_counter() - t0
        self.optimizer.step(grads=self.flat.grad_views())
        n = self.flat.dense_nbytes()
        st.bytes_sent = st.bytes_recv = int(2 * (self.world - 1) / max(1, self.world) * n)
        return st


def build_host_exchange(role: str, flat: FlatModel, cfg, optimizer, group=None) -> Exchange:
    if cfg.exchange == "ps":
        return PSServer(flat, cfg, optimizer, group) if role == "server" else PSWorker(flat, cfg, optimizer, group)
    if cfg.exchange == "allgather":
        return AllgatherExchange(flat, cfg, optimizer, group)
    return AllreduceExchange(flat, cfg, optimizer, group)
