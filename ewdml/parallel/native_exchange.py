"""Exchange strategies on the hand-written NVLink kernels (the product path on B200).

Same ``Exchange`` protocol as :mod:`ewdml.parallel.host_exchange`, but every step is a short chain
of sm_100a kernels on the current CUDA stream and no host collective is called:

* :class:`NativeAllgather` -- K1 (select + QSGD + pack + P2P stores into every peer's mailbox)
  followed by K2 (wait flags, unpack, average, momentum-SGD) -- the "sparse allgather".
* :class:`NativePSWorker` / :class:`NativePSServer` -- K3: workers store payloads into the server
  rank's mailbox; the server decodes/averages/updates and stores weights (method 1/2) or the
  averaged, optionally re-compressed gradient (methods 3/4/5) into every worker's pull mailbox.
* :class:`NativeAllreduce` -- K4 dense fused-scale allreduce in place on the symmetric gradient
  buffer, then the fused SGD kernel.

Mailbox protocol: a mailbox has ``nsrc`` slots x 2 parities and one flag word per slot.  Exchange
number ``q`` (1, 2, ...) uses parity ``q & 1``; the sender's last CTA stores ``q`` into the flag with
release semantics after all payload bytes, the receiver spins with acquire until flag >= q.  Two
parities suffice without acknowledgements: a sender can only reach exchange q+2 after its own
receive of q+1 completed, which needed every peer's send of q+1, which each peer enqueued after its
own receive of q (stream order).
"""
from __future__ import annotations

import ctypes as C
import math
from typing import List, Optional, Sequence, Tuple

import torch

from ..compress.qsgd import code_width
from ..runtime import lib as L
from .flat import FlatModel
from .host_exchange import Exchange, ExchangeStats
from .peer import LocalPeerGroup, PeerGroup, SymmetricBuffer

FLAG_REGION = 4096          # bytes reserved for flags at the start of a mailbox
FLAG_STRIDE = 64            # one flag per 64 B


def make_codec(compressor: str, levels: int = 128, ratio: float = 0.5) -> L.Codec:
    c = L.Codec()
    c.levels, c.topk_ratio = int(levels), float(ratio)
    if compressor == "none":
        c.mode, c.sel, c.val, c.code_bits = L.MODE_DENSE, L.SEL_ALL, L.VAL_RAW32, 32
    elif compressor == "qsgd":
        c.mode, c.sel, c.val, c.code_bits = L.MODE_BITMAP, L.SEL_ALL, L.VAL_QSGD, code_width(levels, True)
    elif compressor == "topk":
        c.mode = L.MODE_IDX16 if ratio <= 1.0 / 16 else L.MODE_BITMAP
        c.sel, c.val, c.code_bits = L.SEL_TOPK, L.VAL_RAW32, 32
    elif compressor == "topk-qsgd":
        c.mode = L.MODE_IDX16 if ratio <= 1.0 / 16 else L.MODE_BITMAP
        c.sel, c.val, c.code_bits = L.SEL_TOPK, L.VAL_QSGD, code_width(levels, True)
    else:
        raise ValueError(compressor)
    return c


class ExchangePlan:
    """Owns an ``ewdml_plan`` (device tables + scratch) for a list of (offset, numel) segments."""

    def __init__(self, offsets: Sequence[int], numels: Sequence[int], codec: L.Codec):
        self.lib = L.load()
        self.codec = codec
        self.nseg = len(offsets)
        self._h = C.c_void_p()
        off = (C.c_int64 * self.nseg)(*offsets)
        num = (C.c_int64 * self.nseg)(*numels)
        L.check(self.lib.ewdml_plan_create(C.byref(self._h), self.nseg, off, num, C.byref(codec)), "plan_create")
        self.slot_bytes = int(self.lib.ewdml_plan_slot_bytes(self._h))
        self.ntiles = int(self.lib.ewdml_plan_num_tiles(self._h))
        self.scratch_bytes = int(self.lib.ewdml_plan_scratch_bytes(self._h))
        lay = (C.c_int64 * 4)()
        L.check(self.lib.ewdml_plan_slot_layout(self._h, lay), "slot_layout")
        self.layout = {"norms": lay[0], "dir": lay[1], "bitmap": lay[2], "data": lay[3]}
        self.total = (offsets[-1] + numels[-1] + 31) // 32 * 32

    @classmethod
    def for_flat(cls, flat: FlatModel, codec: L.Codec) -> "ExchangePlan":
        return cls([s.offset for s in flat.segments], [s.numel for s in flat.segments], codec)

    @staticmethod
    def _stream() -> C.c_void_p:
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def send(self, grad: torch.Tensor, dests: Sequence[Tuple[int, int]], seq: int, seed: int, step: int,
             rng_stream: int, mc_slot: int = 0) -> None:
        d = self._dests(dests, mc_slot)
        L.check(self.lib.ewdml_send(self._h, C.c_void_p(grad.data_ptr()), C.byref(d), seq & 0xFFFFFFFF,
                                    seed & 0xFFFFFFFFFFFFFFFF, step & 0xFFFFFFFF, rng_stream, self._stream()),
                "ewdml_send")

    def recv_apply(self, sources: Sequence[Tuple[int, int]], seq: int, divisor: float, out_mode: int,
                   sgd: Optional[L.Sgd] = None, param: Optional[torch.Tensor] = None,
                   momentum: Optional[torch.Tensor] = None, param_bf16: Optional[torch.Tensor] = None,
                   out: Optional[torch.Tensor] = None) -> None:
        s = L.Sources()
        s.n = len(sources)
        for i, (slot, flag) in enumerate(sources):
            s.slot[i], s.flag[i] = slot, flag
        ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)
        L.check(self.lib.ewdml_recv_apply(self._h, C.byref(s), seq & 0xFFFFFFFF, float(divisor), out_mode,
                                          C.byref(sgd) if sgd is not None else None, ptr(param), ptr(momentum),
                                          ptr(param_bf16), ptr(out), self._stream()), "ewdml_recv_apply")

    @staticmethod
    def _dests(pairs, mc_slot: int = 0) -> L.Dests:
        d = L.Dests()
        d.n = len(pairs)
        for i, (slot, flag) in enumerate(pairs):
            d.slot[i], d.flag[i] = slot, flag
        d.mc_slot = mc_slot or None       # NVLS: one multicast store per vector instead of one store per destination
        return d

    @staticmethod
    def _sources(pairs) -> L.Sources:
        s = L.Sources()
        s.n = len(pairs)
        for i, (slot, flag) in enumerate(pairs):
            s.slot[i], s.flag[i] = slot, flag
        return s

    def send_state(self, grad: torch.Tensor, dests_even, dests_odd, state: torch.Tensor, seed: int, rng_stream: int,
                   mc_even: int = 0, mc_odd: int = 0) -> None:
        """Graph-replayable send: sequence number / RNG step / parity come from the device ``state``.  ``mc_even`` /
        ``mc_odd``: multicast addresses of the slot (NVLS), valid only when the destinations are ALL ranks' copies."""
        L.check(self.lib.ewdml_send_state(self._h, C.c_void_p(grad.data_ptr()), C.byref(self._dests(dests_even, mc_even)),
                                          C.byref(self._dests(dests_odd, mc_odd)), C.c_void_p(state.data_ptr()),
                                          seed & 0xFFFFFFFFFFFFFFFF, rng_stream, self._stream()), "ewdml_send_state")

    def recv_apply_state(self, srcs_even, srcs_odd, state: torch.Tensor, divisor: float, out_mode: int,
                         sgd: Optional[L.Sgd] = None, param=None, momentum=None, param_bf16=None, out=None) -> None:
        ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)
        L.check(self.lib.ewdml_recv_apply_state(self._h, C.byref(self._sources(srcs_even)), C.byref(self._sources(srcs_odd)),
                                                C.c_void_p(state.data_ptr()), float(divisor), out_mode,
                                                C.byref(sgd) if sgd is not None else None, ptr(param), ptr(momentum),
                                                ptr(param_bf16), ptr(out), self._stream()), "ewdml_recv_apply_state")

    def stats(self) -> dict:
        out = (C.c_int64 * 4)()
        L.check(self.lib.ewdml_plan_read_stats(self._h, out, self._stream()), "read_stats")
        return {"entries": out[0], "data_bytes": out[1], "payload_bytes": out[2], "error": out[3]}

    def fallbacks(self) -> int:
        """Segments that took the exact-scan fallback of the top-k selection since this plan was created."""
        out = C.c_int64(0)
        L.check(self.lib.ewdml_plan_fallbacks(self._h, C.byref(out), self._stream()), "plan_fallbacks")
        return int(out.value)

    def seginfo(self):
        out = (C.c_uint32 * (4 * self.nseg))()
        L.check(self.lib.ewdml_plan_read_seginfo(self._h, out, self._stream()), "read_seginfo")
        import struct
        res = []
        for i in range(self.nseg):
            norm = struct.unpack("f", struct.pack("I", out[4 * i + 2]))[0]
            res.append({"thr_key": out[4 * i], "need_eq": out[4 * i + 1], "norm": norm,
                        "ncand": out[4 * i + 3] & 0x7FFFFFFF, "fallback": out[4 * i + 3] >> 31})
        return res

    def close(self):
        if self._h:
            self.lib.ewdml_plan_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Mailbox:
    """Flags + 2 x nsrc payload slots inside one symmetric buffer (same layout on every rank)."""

    def __init__(self, sym: SymmetricBuffer, nsrc: int, slot_bytes: int):
        self.sym, self.nsrc, self.slot_bytes = sym, nsrc, slot_bytes

    @staticmethod
    def nbytes(nsrc: int, slot_bytes: int) -> int:
        return FLAG_REGION + 2 * nsrc * slot_bytes

    def slot(self, owner: int, parity: int, src: int) -> int:
        return self.sym.ptrs[owner] + FLAG_REGION + (parity * self.nsrc + src) * self.slot_bytes

    def flag(self, owner: int, parity: int, src: int) -> int:
        return self.sym.ptrs[owner] + (parity * self.nsrc + src) * FLAG_STRIDE

    def mc_slot(self, parity: int, src: int) -> int:
        """Multicast address of slot (parity, src) in every rank's mailbox, or 0 (no NVLS mapping / switched off)."""
        mc = getattr(self.sym, "mc_ptr", 0)
        return mc + FLAG_REGION + (parity * self.nsrc + src) * self.slot_bytes if mc else 0

    def local_slot_bytes(self, parity: int, src: int) -> torch.Tensor:
        off = FLAG_REGION + (parity * self.nsrc + src) * self.slot_bytes
        return self.sym.local[off:off + self.slot_bytes]


def make_sgd(cfg, first_step: bool, lr_scale: float = 1.0) -> L.Sgd:
    s = L.Sgd()
    s.lr, s.momentum, s.dampening = cfg.lr * lr_scale, cfg.momentum, 0.0
    s.weight_decay, s.nesterov, s.first_step = cfg.weight_decay, int(cfg.nesterov), int(first_step)
    return s


class _NativeBase(Exchange):
    def __init__(self, flat: FlatModel, cfg, optimizer, pg: PeerGroup):
        self.flat, self.cfg, self.optimizer, self.pg = flat, cfg, optimizer, pg
        self.rank, self.world = pg.rank, pg.world
        self.lib = L.load()
        # device step state {seq, step, first, applies}: no per-step value is baked into a launch, so the whole
        # step (including these kernels) can be captured in a CUDA graph and replayed
        self.state = torch.zeros(4, dtype=torch.int32, device=pg.device)
        self._ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        self._last_payload = 0
        from ..config import registry
        self._nvtx = bool(registry.get("nvtx"))            # NVTX ranges around the exchange phases (EWDML_NVTX=1)
        self._poison = bool(registry.get("debug_poison"))  # fill consumed receive slots with 0xFF (EWDML_DEBUG_POISON=1)
        self._host_seq = 0                                 # eager-mode mirror of state[0], for the poison aid only
        if registry.get("trace"):
            self.lib.ewdml_set_trace(1)
        # the fused receive epilogue implements momentum-SGD; with --optimizer adam the kernels deliver the dense averaged
        # gradient instead and a flat whole-model Adam update (tensor ops, graph-capturable) follows
        self.adam = None
        if getattr(cfg, "optimizer", "sgd") == "adam":
            from ..optim.adam import FlatAdam
            self.adam = FlatAdam(flat.param, lr=cfg.lr, weight_decay=cfg.weight_decay, param_bf16=flat.param_bf16)
            self._dense = torch.zeros_like(flat.param)

    def broadcast_initial_weights(self, src: int = 0) -> None:
        if self.world == 1:
            return
        # one dense push of the parameters from `src` through a temporary mailbox
        codec = make_codec("none")
        plan = ExchangePlan.for_flat(self.flat, codec)
        mb = Mailbox(self.pg.alloc("init_weights", Mailbox.nbytes(1, plan.slot_bytes)), 1, plan.slot_bytes)
        if self.rank == src:
            plan.send(self.flat.param, [(mb.slot(r, 1, 0), mb.flag(r, 1, 0)) for r in range(self.world) if r != src],
                      1, 0, 0, 0, mc_slot=mb.mc_slot(1, 0))
        else:
            plan.recv_apply([(mb.slot(self.rank, 1, 0), mb.flag(self.rank, 1, 0))], 1, 1.0, L.OUT_ASSIGN,
                            param=self.flat.param, param_bf16=self.flat.param_bf16)
        torch.cuda.synchronize()
        self.pg.host_barrier("init_weights")
        # BN running statistics start equal as well (they are never exchanged afterwards)
        # -- every rank builds the model from the same seed, so they already are.

    def _rec(self, i: int, stream) -> None:
        if not torch.cuda.is_current_stream_capturing():      # timing events are not recorded into CUDA graphs
            self._ev[i].record(stream)

    def _range(self, name: Optional[str]) -> None:
        """NVTX range boundary: close the open one, open `name` (None = just close)."""
        if not self._nvtx:
            return
        if getattr(self, "_open_range", False):
            torch.cuda.nvtx.range_pop()
            self._open_range = False
        if name is not None:
            torch.cuda.nvtx.range_push("ewdml:" + name)
            self._open_range = True

    def _poison_slots(self, mb: "Mailbox", nsrc: int) -> None:
        """Debug aid: once a slot has been consumed, overwrite it so that a later kernel which reads
        bytes its sender did not write this step decodes garbage instead of silently reusing stale data."""
        if not self._poison:
            return
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("EWDML_DEBUG_POISON is an eager-mode aid; it cannot be captured in a CUDA graph")
        par = self._host_seq & 1
        for r in range(nsrc):
            mb.local_slot_bytes(par, r).fill_(0xFF)

    def _advance(self, bump_seq: bool) -> None:
        self._host_seq += int(bump_seq)
        L.check(self.lib.ewdml_state_advance(C.c_void_p(self.state.data_ptr()), int(bump_seq), ExchangePlan._stream()),
                "state_advance")

    def _sgd_apply(self) -> None:
        if self.adam is not None:
            self.adam.step(self.flat.grad)
            return
        sgd = make_sgd(self.cfg, False)
        f = self.flat
        L.check(self.lib.ewdml_sgd_apply_state(C.c_void_p(f.grad.data_ptr()), C.c_void_p(f.param.data_ptr()),
                                               C.c_void_p(f.momentum.data_ptr()) if f.momentum is not None else None,
                                               C.c_void_p(f.param_bf16.data_ptr()) if f.param_bf16 is not None else None,
                                               f.total, C.byref(sgd), 1.0, C.c_void_p(self.state.data_ptr()),
                                               ExchangePlan._stream()), "sgd_apply")

    def local_step(self) -> ExchangeStats:
        self._advance(False)
        self._sgd_apply()
        return ExchangeStats()

    def _finish_stats(self, st: ExchangeStats, plan, n_remote_out: int, n_remote_in: int,
                      want_stats: bool) -> ExchangeStats:
        if want_stats and plan is not None:
            plans = plan if isinstance(plan, (list, tuple)) else [plan]
            self._last_payload = sum(p.stats()["payload_bytes"] for p in plans)
            self._ev[2].synchronize()
            st.t_send = self._ev[0].elapsed_time(self._ev[1]) * 1e-3
            st.t_recv = self._ev[1].elapsed_time(self._ev[2]) * 1e-3
        st.bytes_sent = n_remote_out * self._last_payload
        st.bytes_recv = n_remote_in * self._last_payload
        return st


class NativeAllgather(_NativeBase):
    """Sparse allgather.  ``--bucket-size`` splits the flat buffer into contiguous buckets (one plan + mailbox each);
    with ``--overlap`` every bucket's K1 + K2 run on a side stream as soon as backward has produced its gradients
    (:mod:`ewdml.parallel.overlap`), otherwise all sends are enqueued first and all receives after them.  Per-tensor
    thresholds / norms and the element-indexed Philox counters make the result independent of the bucketing."""

    def __init__(self, flat, cfg, optimizer, pg: PeerGroup):
        super().__init__(flat, cfg, optimizer, pg)
        from .overlap import partition_buckets
        self.codec = make_codec(cfg.resolved_compressor(), cfg.qsgd_levels, cfg.topk_ratio)
        self.buckets = partition_buckets(flat.segments, cfg.bucket_size)
        self.plans, self.mbs = [], []
        for b, (lo, hi) in enumerate(self.buckets):
            segs = flat.segments[lo:hi]
            plan = ExchangePlan([s.offset for s in segs], [s.numel for s in segs], self.codec)
            name = "allgather" if len(self.buckets) == 1 else f"allgather_b{b}"
            self.plans.append(plan)
            self.mbs.append(Mailbox(pg.alloc(name, Mailbox.nbytes(self.world, plan.slot_bytes)), self.world, plan.slot_bytes))
        self.plan, self.mb = self.plans[0], self.mbs[0]
        self.side: Optional[torch.cuda.Stream] = None
        # error feedback (optional; the reference has none): what compression dropped is added to the next gradient
        self.residual = torch.zeros_like(flat.grad) if cfg.error_feedback else None
        self._own = torch.zeros_like(flat.grad) if cfg.error_feedback else None

    def _send(self, b: int) -> None:
        me, W, mb = self.rank, self.world, self.mbs[b]
        src = self.flat.grad if self.residual is None else self.residual
        self.plans[b].send_state(src, [(mb.slot(r, 0, me), mb.flag(r, 0, me)) for r in range(W)],
                                 [(mb.slot(r, 1, me), mb.flag(r, 1, me)) for r in range(W)], self.state, self.cfg.seed, me,
                                 mb.mc_slot(0, me), mb.mc_slot(1, me))     # all W ranks receive: one multicast store each

    def _recv(self, b: int) -> None:
        me, W, mb = self.rank, self.world, self.mbs[b]
        if self.residual is not None:
            # decode my own payload (already in my mailbox) and keep what it lost: residual = (grad + residual) - sent
            self.plans[b].recv_apply_state([(mb.slot(me, 0, me), mb.flag(me, 0, me))], [(mb.slot(me, 1, me), mb.flag(me, 1, me))],
                                           self.state, 1.0, L.OUT_GRAD, out=self._own)
            lo, hi = self.buckets[b]
            o0, o1 = self.flat.segments[lo].offset, self.flat.segments[hi - 1].offset + self.flat.segments[hi - 1].numel
            self.residual[o0:o1].sub_(self._own[o0:o1])
        srcs = ([(mb.slot(me, 0, r), mb.flag(me, 0, r)) for r in range(W)], [(mb.slot(me, 1, r), mb.flag(me, 1, r)) for r in range(W)])
        if self.adam is not None:       # dense average of this bucket's tensors; Adam runs once after the last bucket
            self.plans[b].recv_apply_state(srcs[0], srcs[1], self.state, float(W), L.OUT_GRAD, out=self._dense)
            return
        self.plans[b].recv_apply_state(srcs[0], srcs[1], self.state, float(W),
                                       L.OUT_SGD, make_sgd(self.cfg, False), self.flat.param, self.flat.momentum,
                                       self.flat.param_bf16)

    def _fold_residual(self) -> None:
        if self.residual is not None:
            self.residual.add_(self.flat.grad)

    def step(self, step_idx: int, is_sync_step: bool = True, want_stats: bool = False) -> ExchangeStats:
        if not is_sync_step:
            return self.local_step()
        W = self.world
        cur = torch.cuda.current_stream()
        self._advance(True)
        self._fold_residual()
        self._rec(0, cur)
        self._range("select+encode+push")
        for b in range(len(self.plans) - 1, -1, -1):
            self._send(b)
        self._rec(1, cur)
        self._range("wait+decode+apply")
        for b in range(len(self.plans) - 1, -1, -1):
            self._recv(b)
            self._poison_slots(self.mbs[b], W)
        if self.adam is not None:
            self.adam.step(self._dense)
        self._range(None)
        self._rec(2, cur)
        return self._finish_stats(ExchangeStats(), self.plans, W - 1, W - 1, want_stats)

    # ---- backward-overlapped form: begin_overlap() before forward, bucket_ready(b) from the scheduler, finish after ----
    def begin_overlap(self) -> None:
        if self.residual is not None:
            raise RuntimeError("--overlap and --error-feedback cannot be combined (the residual is folded in after backward)")
        cur = torch.cuda.current_stream()
        if self.side is None:
            self.side = torch.cuda.Stream(device=self.pg.device, priority=-1)
        self._advance(True)
        ev = torch.cuda.Event()
        ev.record(cur)
        self.side.wait_event(ev)

    def bucket_ready(self, b: int) -> None:
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self.side.wait_event(ev)
        with torch.cuda.stream(self.side):
            self._send(b)
            self._recv(b)

    def finish_overlap(self, want_stats: bool = False) -> ExchangeStats:
        torch.cuda.current_stream().wait_stream(self.side)
        if self.adam is not None:
            self.adam.step(self._dense)
        W = self.world
        if want_stats:      # payload size only: the phases overlap backward, so there is no separate send / receive time
            self._last_payload = sum(p.stats()["payload_bytes"] for p in self.plans)
        return self._finish_stats(ExchangeStats(), self.plans, W - 1, W - 1, False)


class NativePSWorker(_NativeBase):
    def __init__(self, flat, cfg, optimizer, pg: PeerGroup, role: str = "worker"):
        super().__init__(flat, cfg, optimizer, pg)
        W = self.world
        self.role = role
        self.num_workers = W - 1
        self.push_codec = make_codec(cfg.resolved_compressor(), cfg.qsgd_levels, cfg.topk_ratio)
        self.push = ExchangePlan.for_flat(flat, self.push_codec)
        self.pull_mode = cfg.resolved_pull()
        pull_comp = cfg.resolved_compressor() if (cfg.resolved_compress_pull() and self.pull_mode == "grads") else "none"
        self.pull_codec = make_codec(pull_comp, cfg.qsgd_levels, cfg.topk_ratio)
        self.pull = ExchangePlan.for_flat(flat, self.pull_codec)
        self.push_mb = Mailbox(pg.alloc("ps_push", Mailbox.nbytes(W, self.push.slot_bytes)), W, self.push.slot_bytes)
        self.pull_mb = Mailbox(pg.alloc("ps_pull", Mailbox.nbytes(1, self.pull.slot_bytes)), 1, self.pull.slot_bytes)
        self.avg = torch.zeros(flat.total, dtype=torch.float32, device=flat.param.device) if role == "server" else None

    def step(self, step_idx: int, is_sync_step: bool = True, want_stats: bool = False) -> ExchangeStats:
        if not is_sync_step:
            return self.local_step() if self.role == "worker" else ExchangeStats()
        me, W = self.rank, self.world
        push, pull = self.push_mb, self.pull_mb
        cur = torch.cuda.current_stream()
        st = ExchangeStats()
        self._advance(True)
        self._rec(0, cur)
        sgd = make_sgd(self.cfg, False)
        if self.role == "worker":
            self.push.send_state(self.flat.grad, [(push.slot(0, 0, me), push.flag(0, 0, me))],
                                 [(push.slot(0, 1, me), push.flag(0, 1, me))], self.state, self.cfg.seed, me)
            self._rec(1, cur)
            src = [[(pull.slot(me, par, 0), pull.flag(me, par, 0))] for par in (0, 1)]
            if self.pull_mode == "weights":
                self.pull.recv_apply_state(src[0], src[1], self.state, 1.0, L.OUT_ASSIGN, param=self.flat.param,
                                           param_bf16=self.flat.param_bf16)
            elif self.adam is not None:
                self.pull.recv_apply_state(src[0], src[1], self.state, 1.0, L.OUT_GRAD, out=self._dense)
                self.adam.step(self._dense)
            else:
                self.pull.recv_apply_state(src[0], src[1], self.state, 1.0, L.OUT_SGD, sgd, self.flat.param,
                                           self.flat.momentum, self.flat.param_bf16)
            self._rec(2, cur)
            return self._finish_stats(st, self.push, 1, 1, want_stats)
        # ---- server ------------------------------------------------------------------------
        srcs = [[(push.slot(0, par, r), push.flag(0, par, r)) for r in range(1, W)] for par in (0, 1)]
        dests = [[(pull.slot(r, par, 0), pull.flag(r, par, 0)) for r in range(1, W)] for par in (0, 1)]
        if self.pull_mode == "weights":
            if self.adam is not None:
                self.push.recv_apply_state(srcs[0], srcs[1], self.state, float(self.num_workers), L.OUT_GRAD, out=self._dense)
                self.adam.step(self._dense)
            else:
                self.push.recv_apply_state(srcs[0], srcs[1], self.state, float(self.num_workers), L.OUT_SGD, sgd,
                                           self.flat.param, self.flat.momentum, self.flat.param_bf16)
            self._rec(1, cur)
            # weight pull = ONE multicast store per vector (the server's own unused pull slot receives a copy too)
            self.pull.send_state(self.flat.param, dests[0], dests[1], self.state, self.cfg.seed, 0,
                                 pull.mc_slot(0, 0), pull.mc_slot(1, 0))
        else:
            self.push.recv_apply_state(srcs[0], srcs[1], self.state, float(self.num_workers), L.OUT_GRAD, out=self.avg)
            self._rec(1, cur)
            self.pull.send_state(self.avg, dests[0], dests[1], self.state, self.cfg.seed, 0,
                                 pull.mc_slot(0, 0), pull.mc_slot(1, 0))
        self._rec(2, cur)
        return self._finish_stats(st, self.pull, self.num_workers, self.num_workers, want_stats)


class NativePSServer(NativePSWorker):
    def __init__(self, flat, cfg, optimizer, pg: PeerGroup):
        super().__init__(flat, cfg, optimizer, pg, role="server")


class NativeAllreduce(_NativeBase):
    """Needs ``flat.grad`` to live in a symmetric buffer of >= 2*total floats (see native_allocator)."""

    def __init__(self, flat, cfg, optimizer, pg: PeerGroup, grad_sym: SymmetricBuffer, predivide: float = 1.0):
        super().__init__(flat, cfg, optimizer, pg)
        self.grad_sym = grad_sym
        self.flags = pg.alloc("ar_flags", 4096)
        self.comm = L.Comm()
        self.comm.rank, self.comm.world = self.rank, self.world
        for r in range(self.world):
            self.comm.bufs[r] = grad_sym.ptrs[r]
            self.comm.peer_flags[r] = self.flags.ptrs[r]
        self.comm.flags = self.flags.ptrs[self.rank]
        self.comm.mc_buf = getattr(grad_sym, "mc_ptr", 0) or None      # NVLS: in-switch reduction + multicast store
        self.work_counter = self.flags.ptrs[self.rank] + 2048
        from ..config import registry
        self.algo = {"auto": L.AR_AUTO, "oneshot": L.AR_ONESHOT, "twoshot": L.AR_TWOSHOT,
                     "nvls": L.AR_NVLS}[registry.get("allreduce_algo")]

    def allreduce_(self, n: int, scale: float) -> None:
        L.check(self.lib.ewdml_allreduce_state(C.byref(self.comm), n, scale, self.algo, 0, C.c_void_p(self.state.data_ptr()),
                                               C.c_void_p(self.work_counter), ExchangePlan._stream()), "allreduce")

    def step(self, step_idx: int, is_sync_step: bool = True, want_stats: bool = False) -> ExchangeStats:
        if not is_sync_step:
            return self.local_step()
        cur = torch.cuda.current_stream()
        self._advance(True)
        self._rec(0, cur)
        self.allreduce_(self.flat.total, 1.0 / self.world)
        self._rec(1, cur)
        self._sgd_apply()
        st = ExchangeStats()
        self._rec(2, cur)
        n = self.flat.dense_nbytes()
        st.bytes_sent = st.bytes_recv = int((self.world - 1) / self.world * n)
        if want_stats:
            self._ev[2].synchronize()
            st.t_send = self._ev[0].elapsed_time(self._ev[1]) * 1e-3
            st.t_recv = self._ev[1].elapsed_time(self._ev[2]) * 1e-3
        return st


def native_allocator(pg: PeerGroup, total_hint: dict):
    """FlatModel allocator that places the gradient buffer in the symmetric heap (2x for one-shot)."""
    def alloc(tag: str, n: int, dtype: torch.dtype) -> torch.Tensor:
        if tag == "grad":
            sym = pg.alloc("flat_grad", 2 * n * 4)
            total_hint["grad_sym"] = sym
            return sym.view(torch.float32, 0, n)
        return torch.zeros(n, dtype=dtype, device=pg.device)
    return alloc


def make_native_factory(device: torch.device, pg: Optional[PeerGroup] = None):
    """Factory with the signature of ``build_host_exchange`` for the CLI / trainers."""
    state = {"pg": pg}

    def factory(role: str, flat: FlatModel, cfg, optimizer, group=None):
        if state["pg"] is None:
            state["pg"] = PeerGroup.from_env(device)
        g = state["pg"]
        if cfg.exchange == "ps":
            return NativePSServer(flat, cfg, optimizer, g) if role == "server" else NativePSWorker(flat, cfg, optimizer, g)
        if cfg.exchange == "allgather":
            return NativeAllgather(flat, cfg, optimizer, g)
        sym = getattr(flat, "grad_sym", None)
        if sym is None:
            raise RuntimeError("--exchange allreduce on the native backend needs FlatModel(allocator=native_allocator(...))")
        return NativeAllreduce(flat, cfg, optimizer, g, sym)

    factory.peer_group = lambda: state["pg"]
    return factory
