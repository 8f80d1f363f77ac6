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
# Parameter server with backup workers / straggler cut-off
# --------------------------------------------------------------------------------------
_PUSH_TAG, _PULL_TAG, _HDR = 11, 12, 16


class PSWorkerP2P(PSWorker):
    """Worker side of the partial-aggregation protocol: point-to-point ``send`` of (step header + payload) and a
    ``recv`` of the pull message instead of the ``gather`` / ``broadcast`` pair, so the server is free to close a
    step without this worker."""

    def __init__(self, flat, cfg, optimizer, group=None):
        super().__init__(flat, cfg, optimizer, group)
        self._msg = torch.zeros(_HDR + self.push.nbytes, dtype=torch.uint8)

    def step(self, step_idx: int, is_sync_step: bool = True) -> ExchangeStats:
        if not is_sync_step:
            return self.local_step()
        st = ExchangeStats()
        t0 = time.perf_counter()
        self._msg[:8].view(torch.int64)[0] = step_idx
        self.push.encode(self._ef_source(), step_idx, out=self._msg[_HDR:])
        self._ef_update(self.push, self._msg[_HDR:])
        dist.send(self._msg, dst=0, group=self.group, tag=_PUSH_TAG)
        st.t_send = time.perf_counter() - t0
        st.bytes_sent = self._msg.numel()
        t0 = time.perf_counter()
        if self.pull_mode == "weights":
            dist.recv(self._wbuf, src=0, group=self.group, tag=_PULL_TAG)
            self.flat.load_flat_params(self._wbuf.to(self.flat.param.device))
            st.bytes_recv = self._wbuf.numel() * 4
        else:
            dist.recv(self._pull_msg, src=0, group=self.group, tag=_PULL_TAG)
            self.pull.decode(self._pull_msg, out=self._dense)
            st.bytes_recv = self._pull_msg.numel()
            dense = self._dense.to(self.flat.param.device)
            self.optimizer.step(grads=[self.flat.view(dense, i) for i in range(len(self.flat.segments))])
        st.t_recv = time.perf_counter() - t0
        return st


class PSServerP2P(PSServer):
    """Aggregates the first ``num_aggregate`` gradients of every step (backup workers, Chen et al. 2016 -- the
    semantics the reference's ``--num-aggregate`` flag names but never implements, ``distributed_nn.py:50-59``); with
    ``--mode kill`` it additionally stops waiting ``kill_threshold`` seconds after the first arrival (the dead
    straggler-kill loop of ``model_ops/lenet.py:188-255``; here the straggler is not interrupted, its late gradient
    is dropped).  A receiver thread takes pushes from any source in arrival order and queues them by step; gradients
    of steps that are already closed are counted in ``dropped``.  ``--timeout-s`` bounds the wait for a dead worker."""

    def __init__(self, flat, cfg, optimizer, group=None):
        import queue
        import threading
        super().__init__(flat, cfg, optimizer, group)
        self.need = max(1, min(cfg.num_aggregate, self.num_workers))
        self.kill = cfg.mode == "kill"
        self.threshold = float(cfg.kill_threshold)
        self.timeout = float(cfg.timeout_s)
        self.dropped = 0
        self.used_per_step: List[int] = []
        self._q: "queue.Queue" = queue.Queue()
        self._stop = False
        self._received = 0
        self._out: list = []
        self._nbytes = _HDR + self.decoders[0].nbytes
        self._thread = threading.Thread(target=self._receiver, daemon=True)
        self._expected: Optional[int] = None        # set by expect(): total number of pushes the job will produce
        self._started = False

    def expect(self, sync_steps: int) -> None:
        self._expected = sync_steps * self.num_workers

    def _receiver(self) -> None:
        while not self._stop and (self._expected is None or self._received < self._expected):
            buf = torch.empty(self._nbytes, dtype=torch.uint8)
            try:
                src = dist.recv(buf, group=self.group, tag=_PUSH_TAG)          # any source, arrival order
            except Exception as e:                                             # group torn down / peer died
                self._q.put(("error", e))
                return
            self._received += 1
            self._q.put((int(src), int(buf[:8].view(torch.int64)[0]), buf))

    def step(self, step_idx: int, is_sync_step: bool = True) -> ExchangeStats:
        import queue
        if not is_sync_step:
            return ExchangeStats()
        if not self._started:
            self._thread.start()
            self._started = True
        st = ExchangeStats()
        t0 = time.perf_counter()
        self._out = [(w, b) for (w, b) in self._out if not _done(w)]
        while len(self._out) > 8 * self.num_workers:        # back-pressure: at most 8 pulls in flight per worker
            self._out.pop(0)[0].wait()
        got, first = [], None
        while len(got) < self.need:
            now = time.perf_counter()
            if self.kill and first is not None:
                wait = self.threshold - (now - first)
                if wait <= 0:
                    break
            else:
                wait = self.timeout - (now - t0)
                if wait <= 0:
                    raise TimeoutError(f"parameter server: {len(got)} of {self.num_workers} gradients after {self.timeout}s "
                                       f"at step {step_idx} (a worker died?)")
            try:
                item = self._q.get(timeout=max(wait, 1e-3))
            except queue.Empty:
                continue
            if item[0] == "error":
                raise RuntimeError(f"parameter server receiver failed: {item[1]}")
            src, step, buf = item
            if step != step_idx:
                self.dropped += 1                   # a straggler's gradient for a step that is already closed
                continue
            got.append((src, buf))
            first = first or time.perf_counter()
        self._avg.zero_()
        for src, buf in sorted(got, key=lambda t: t[0]):
            self.decoders[src].decode(buf[_HDR:], out=self._avg, accumulate=True)
        self._avg.div_(len(got))
        self.used_per_step.append(len(got))
        st.t_recv = time.perf_counter() - t0
        st.bytes_recv = len(got) * self._nbytes
        t0 = time.perf_counter()
        if self.pull_mode == "weights":
            avg = self._avg.to(self.flat.param.device)
            self.optimizer.step(grads=[self.flat.view(avg, i) for i in range(len(self.flat.segments))])
            msg = self.flat.param.detach().cpu().clone()
            st.bytes_sent = self.num_workers * msg.numel() * 4
        else:
            msg = self.pull.encode(self._avg, step_idx).clone()
            st.bytes_sent = self.num_workers * msg.numel()
        for r in range(1, self.world):
            self._out.append((dist.isend(msg, dst=r, group=self.group, tag=_PULL_TAG), msg))
        st.t_send = time.perf_counter() - t0
        return st

    def drain(self) -> None:
        """End of training: let the receiver take (and drop) what stragglers still owe, finish the outstanding sends."""
        for w, _ in self._out:
            w.wait()
        self._out.clear()
        if self._started and self._expected is not None:
            self._thread.join(timeout=self.timeout)
        self._stop = True
        while not self._q.empty():
            item = self._q.get_nowait()
            if item[0] != "error":
                self.dropped += 1


class PSServerAsync(PSServerP2P):
    """Asynchronous parameter server (``--mode async``): the variant the report describes next to the synchronous one
    (``Report.zip!main.tex:58-59``) but the reference never implements.  Each gradient is applied the moment it
    arrives and the fresh weights go back to that worker only; workers never wait for each other, so a slow worker
    costs nobody else anything -- at the price of stale gradients.  One call of ``step`` serves one arrival; the
    training loop calls it ``workers x steps`` times."""

    def __init__(self, flat, cfg, optimizer, group=None):
        super().__init__(flat, cfg, optimizer, group)
        self.staleness: List[int] = []
        self._version = 0
        self._seen = {r: 0 for r in range(1, self.world)}

    def step(self, step_idx: int, is_sync_step: bool = True) -> ExchangeStats:
        import queue
        if not self._started:
            self._thread.start()
            self._started = True
        st = ExchangeStats()
        t0 = time.perf_counter()
        try:
            item = self._q.get(timeout=self.timeout)
        except queue.Empty:
            raise TimeoutError(f"asynchronous parameter server: no gradient for {self.timeout}s (workers died?)")
        if item[0] == "error":
            raise RuntimeError(f"parameter server receiver failed: {item[1]}")
        src, wstep, buf = item
        self.decoders[src].decode(buf[_HDR:], out=self._avg, accumulate=False)
        st.t_recv = time.perf_counter() - t0
        st.bytes_recv = self._nbytes
        t0 = time.perf_counter()
        avg = self._avg.to(self.flat.param.device)
        self.optimizer.step(grads=[self.flat.view(avg, i) for i in range(len(self.flat.segments))])
        self.staleness.append(self._version - self._seen[src])      # updates applied since this worker last pulled
        self._version += 1
        self._seen[src] = self._version
        msg = self.flat.param.detach().cpu().clone()
        self._out.append((dist.isend(msg, dst=src, group=self.group, tag=_PULL_TAG), msg))
        while len(self._out) > 8 * self.num_workers:
            self._out.pop(0)[0].wait()
        st.bytes_sent = msg.numel() * 4
        st.t_send = time.perf_counter() - t0
        self.used_per_step.append(1)
        return st


class PSWorkerAsync(PSWorkerP2P):
    """Pushes its gradient, pulls whatever the weights are by then (always a weights pull)."""

    def __init__(self, flat, cfg, optimizer, group=None):
        super().__init__(flat, cfg, optimizer, group)
        self.pull_mode = "weights"


def _done(work) -> bool:
    try:
        return bool(work.is_completed())
    except Exception:
        return False


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
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        if cfg.mode == "async":
            return PSServerAsync(flat, cfg, optimizer, group) if role == "server" else PSWorkerAsync(flat, cfg, optimizer, group)
        if cfg.mode == "kill" or cfg.num_aggregate < world - 1:     # backup workers / straggler cut-off
            return PSServerP2P(flat, cfg, optimizer, group) if role == "server" else PSWorkerP2P(flat, cfg, optimizer, group)
        return PSServer(flat, cfg, optimizer, group) if role == "server" else PSWorker(flat, cfg, optimizer, group)
    if cfg.exchange == "allgather":
        return AllgatherExchange(flat, cfg, optimizer, group)
    return AllreduceExchange(flat, cfg, optimizer, group)
