"""Collective engine on tensors: the Python face of ``csrc/comm/collectives.cu`` (work-list row N3).

This is what a user of the reference's communication layer reaches for when they need a plain
collective on a GPU tensor -- the ``MPI_Allreduce / Allgather / Bcast / Gather / Scatter / Alltoall /
Reduce / Reduce_scatter / Barrier / Send / Recv`` surface that Horovod and mpi4py would have driven
through the vendored Open MPI (``ompi:ompi/mpi/c/allreduce.c:46-117`` dispatching through
``comm->c_coll``, ``ompi:ompi/mca/coll/coll.h:142-255``; gloo ``gather``/``broadcast`` in
``src/distributed_worker.py:256,278,350``).  Here every operation is ONE sm_100a kernel that stores into
the destination ranks' symmetric windows over NVLink and then consumes the local window; there is no
host staging (cf. ``ompi:ompi/mca/coll/cuda/coll_cuda_allreduce.c:30-74``) and no ring or tree.

* blocking form: enqueued on the current CUDA stream (stream-ordered like any torch op);
* non-blocking form (``coll/libnbc``'s role, ``ompi:ompi/mca/coll/libnbc/nbc_iallreduce.c:227``):
  ``i*`` methods run the same kernel on the engine's side stream and return a :class:`Request`
  whose ``wait()`` makes the current stream wait for it;
* decision layer (``coll/tuned``): :meth:`select_algo` / :meth:`set_algo`, plus the environment
  variables ``EWDML_COLL_ALLREDUCE_ALGO`` and ``EWDML_COLL_ONESHOT_MAX_BYTES``.

The training exchanges (K1-K4) do not go through this engine: they are fused with their producer /
consumer math.  The engine serves everything else: metric averaging, parameter / optimizer-state
broadcast (``hvd.broadcast_parameters``), BatchNorm statistics, user code.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch

from ..runtime import lib as L
from .peer import LocalPeerGroup, PeerGroup, SymmetricBuffer

OPS = {"bcast": L.COLL_BCAST, "gather": L.COLL_GATHER, "scatter": L.COLL_SCATTER, "allgather": L.COLL_ALLGATHER,
       "alltoall": L.COLL_ALLTOALL, "reduce": L.COLL_REDUCE, "reduce_scatter": L.COLL_REDUCE_SCATTER,
       "allreduce": L.COLL_ALLREDUCE, "send": L.COLL_SEND, "recv": L.COLL_RECV, "barrier": L.COLL_BARRIER}
ALGOS = {"auto": L.ALGO_AUTO, "oneshot": L.ALGO_ONESHOT, "twoshot": L.ALGO_TWOSHOT}


class Request:
    """Handle of a non-blocking collective (a CUDA event on the engine's side stream)."""

    def __init__(self, event: torch.cuda.Event, keep=None):
        self._event, self._keep = event, keep

    def wait(self) -> None:
        torch.cuda.current_stream().wait_event(self._event)

    def synchronize(self) -> None:
        self._event.synchronize()

    def test(self) -> bool:
        return self._event.query()


def _nbytes(t: torch.Tensor) -> int:
    return t.numel() * t.element_size()


class CollectiveEngine:
    def __init__(self, rank: int, world: int, win: SymmetricBuffer, flags: SymmetricBuffer, slot_payload: int,
                 device: torch.device):
        self.lib = L.load()
        self.rank, self.world, self.device = rank, world, device
        self._win, self._flags = win, flags          # keep the allocations alive
        self._h = C.c_void_p()
        wins = (C.c_void_p * world)(*win.ptrs[:world])
        fl = (C.c_void_p * world)(*flags.ptrs[:world])
        L.check(self.lib.ewdml_engine_create(C.byref(self._h), rank, world, wins, fl, slot_payload), "engine_create")
        self.slot_payload = int(self.lib.ewdml_engine_slot_payload(self._h))
        self._side: Optional[torch.cuda.Stream] = None
        self._last_async: Optional[torch.cuda.Event] = None     # last non-blocking op still possibly in flight

    # ---- construction ------------------------------------------------------------------------------
    @staticmethod
    def _sizes(world: int, slot_payload: int):
        lib = L.load()
        slot_payload = max(256, (slot_payload + 255) // 256 * 256)
        return slot_payload, int(lib.ewdml_engine_window_bytes(world, slot_payload)), int(lib.ewdml_engine_flag_bytes())

    @classmethod
    def create(cls, pg: PeerGroup, slot_payload: int = 4 << 20, name: str = "coll", heap=None) -> "CollectiveEngine":
        """Collective over the whole peer group (every rank calls this at the same point).  With ``heap`` (a
        :class:`ewdml.parallel.peer.SymmetricHeap`) the window and flags are carved out of the already-mapped region,
        so creating an engine costs no IPC exchange."""
        sp, wbytes, fbytes = cls._sizes(pg.world, slot_payload)
        if heap is not None:
            win, flags = heap.alloc(name + "_win", wbytes), heap.alloc(name + "_flags", fbytes)
            flags.local.zero_()                       # recycled heap memory: flags must start at zero on every rank
            torch.cuda.synchronize(pg.device)
            pg.host_barrier(name + "_ready")
        else:
            win = pg.alloc(name + "_win", wbytes)
            flags = pg.alloc(name + "_flags", fbytes)
        return cls(pg.rank, pg.world, win, flags, sp, pg.device)

    @classmethod
    def create_local(cls, lpg: LocalPeerGroup, slot_payload: int = 1 << 20, name: str = "coll") -> List["CollectiveEngine"]:
        """One engine per emulated rank on one GPU (tests): drive them with PHASE_PUT on every rank, then PHASE_WAIT."""
        sp, wbytes, fbytes = cls._sizes(lpg.world, slot_payload)
        wins = lpg.alloc(name + "_win", wbytes)
        flags = lpg.alloc(name + "_flags", fbytes)
        return [cls(r, lpg.world, wins[r], flags[r], sp, lpg.device) for r in range(lpg.world)]

    def close(self) -> None:
        if self._h:
            self.lib.ewdml_engine_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- decision layer ----------------------------------------------------------------------------
    def set_algo(self, op: str, algo: str) -> None:
        L.check(self.lib.ewdml_engine_set_algo(self._h, OPS[op], ALGOS[algo]), "engine_set_algo")

    def select_algo(self, op: str, nbytes: int) -> str:
        a = int(self.lib.ewdml_engine_select_algo(self._h, OPS[op], int(nbytes)))
        return {v: k for k, v in ALGOS.items()}[a]

    def calls(self, op: str) -> int:
        return int(self.lib.ewdml_engine_calls(self._h, OPS[op]))

    # ---- plumbing -----------------------------------------------------------------------------------
    def _stream(self) -> C.c_void_p:
        """Stream handle for the next engine kernel.  The engine has ONE window, one set of DATA/ACK flags and one pair of
        CTA-election counters, so only one of its kernels may run at a time: a call issued on the current stream first
        waits for the last non-blocking operation (which runs on the side stream), and the side stream in turn waits for
        the current stream before every ``i*`` call -- engine work is therefore totally ordered whatever mix of blocking
        and non-blocking calls the user makes (``r = iall_reduce(a); all_reduce(b); r.wait()`` is safe)."""
        cur = torch.cuda.current_stream()
        if self._last_async is not None and (self._side is None or cur != self._side):
            cur.wait_event(self._last_async)
        return C.c_void_p(cur.cuda_stream)

    def _check(self, t: Optional[torch.Tensor], what: str) -> None:
        if t is None:
            return
        if not t.is_cuda or not t.is_contiguous():
            raise ValueError(f"{what}: need a contiguous CUDA tensor")
        if _nbytes(t) % 16 or t.data_ptr() % 16:
            raise ValueError(f"{what}: size and address must be multiples of 16 bytes (got {_nbytes(t)} B); "
                             "use the padded helpers broadcast()/all_reduce() or pad the tensor")

    def _run(self, op: int, src: Optional[torch.Tensor], dst: Optional[torch.Tensor], n: int, root: int = 0,
             scale: float = 1.0, phases: int = L.PHASE_BOTH) -> None:
        p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)
        L.check(self.lib.ewdml_coll(self._h, op, p(src), p(dst), int(n), int(root), float(scale), int(phases), self._stream()),
                "ewdml_coll(" + self.lib.ewdml_coll_op_name(op).decode() + ")")

    def _padded(self, t: torch.Tensor):
        """(work tensor, needs copy back) with a 16-byte multiple size."""
        nb = _nbytes(t)
        if t.is_contiguous() and nb % 16 == 0 and t.data_ptr() % 16 == 0:
            return t, False
        pad = torch.zeros((nb + 15) // 16 * 16 // t.element_size(), dtype=t.dtype, device=t.device)
        pad[:t.numel()].copy_(t.reshape(-1))
        return pad, True

    # ---- blocking (stream-ordered) collectives ------------------------------------------------------
    def broadcast(self, t: torch.Tensor, root: int = 0, phases: int = L.PHASE_BOTH) -> torch.Tensor:
        w, back = self._padded(t)
        self._run(L.COLL_BCAST, None, w, _nbytes(w), root, 1.0, phases)
        if back and (phases & L.PHASE_WAIT):
            t.copy_(w[:t.numel()].view_as(t))
        return t

    def gather(self, src: torch.Tensor, dst: Optional[torch.Tensor], root: int = 0, phases: int = L.PHASE_BOTH):
        self._check(src, "gather src"); self._check(dst, "gather dst")
        if self.rank == root and (dst is None or _nbytes(dst) != self.world * _nbytes(src)):
            raise ValueError("gather: the root needs dst of world * src bytes")
        self._run(L.COLL_GATHER, src, dst if self.rank == root else src, _nbytes(src), root, 1.0, phases)
        return dst

    def scatter(self, src: Optional[torch.Tensor], dst: torch.Tensor, root: int = 0, phases: int = L.PHASE_BOTH):
        self._check(src, "scatter src"); self._check(dst, "scatter dst")
        if self.rank == root and (src is None or _nbytes(src) != self.world * _nbytes(dst)):
            raise ValueError("scatter: the root needs src of world * dst bytes")
        self._run(L.COLL_SCATTER, src if self.rank == root else dst, dst, _nbytes(dst), root, 1.0, phases)
        return dst

    def all_gather(self, src: torch.Tensor, dst: torch.Tensor, phases: int = L.PHASE_BOTH) -> torch.Tensor:
        self._check(src, "all_gather src"); self._check(dst, "all_gather dst")
        if _nbytes(dst) != self.world * _nbytes(src):
            raise ValueError("all_gather: dst must hold world * src bytes")
        self._run(L.COLL_ALLGATHER, src, dst, _nbytes(src), 0, 1.0, phases)
        return dst

    def all_to_all(self, src: torch.Tensor, dst: torch.Tensor, phases: int = L.PHASE_BOTH) -> torch.Tensor:
        self._check(src, "all_to_all src"); self._check(dst, "all_to_all dst")
        if _nbytes(src) != _nbytes(dst) or _nbytes(src) % (16 * self.world):
            raise ValueError("all_to_all: src and dst hold world equal pieces of a multiple of 16 bytes")
        self._run(L.COLL_ALLTOALL, src, dst, _nbytes(src) // self.world, 0, 1.0, phases)
        return dst

    def reduce(self, src: torch.Tensor, dst: Optional[torch.Tensor], root: int = 0, scale: float = 1.0,
               phases: int = L.PHASE_BOTH):
        self._check(src, "reduce src"); self._check(dst, "reduce dst")
        self._f32(src)
        if self.rank == root and dst is None:
            raise ValueError("reduce: the root needs dst")
        self._run(L.COLL_REDUCE, src, dst if self.rank == root else src, _nbytes(src), root, scale, phases)
        return dst

    def reduce_scatter(self, src: torch.Tensor, dst: torch.Tensor, scale: float = 1.0, phases: int = L.PHASE_BOTH):
        self._check(src, "reduce_scatter src"); self._check(dst, "reduce_scatter dst")
        self._f32(src); self._f32(dst)
        if _nbytes(src) != self.world * _nbytes(dst):
            raise ValueError("reduce_scatter: src must hold world * dst bytes")
        self._run(L.COLL_REDUCE_SCATTER, src, dst, _nbytes(dst), 0, scale, phases)
        return dst

    def all_reduce(self, t: torch.Tensor, out: Optional[torch.Tensor] = None, scale: float = 1.0,
                   phases: int = L.PHASE_BOTH) -> torch.Tensor:
        """fp32 sum over ranks (in rank order, so every rank gets bit-identical results) times ``scale``."""
        self._f32(t)
        out = t if out is None else out
        w, back = self._padded(t)
        o = w if (back or out is t) else out
        if o is not w:
            self._check(o, "all_reduce out")
        self._run(L.COLL_ALLREDUCE, w, o, _nbytes(w), 0, scale, phases)
        if back and (phases & L.PHASE_WAIT):
            out.copy_(w[:t.numel()].view_as(out))
        return out

    def barrier(self, phases: int = L.PHASE_BOTH) -> None:
        self._run(L.COLL_BARRIER, None, None, 0, 0, 1.0, phases)

    def send(self, t: torch.Tensor, peer: int, tag: int = 0) -> None:
        self._check(t, "send")
        L.check(self.lib.ewdml_coll_send(self._h, C.c_void_p(t.data_ptr()), _nbytes(t), peer, tag, self._stream()), "coll_send")

    def recv(self, t: torch.Tensor, peer: int, tag: int = 0) -> torch.Tensor:
        self._check(t, "recv")
        L.check(self.lib.ewdml_coll_recv(self._h, C.c_void_p(t.data_ptr()), _nbytes(t), peer, tag, self._stream()), "coll_recv")
        return t

    def sendrecv(self, src: torch.Tensor, to: int, dst: torch.Tensor, frm: int, tag: int = 0) -> torch.Tensor:
        """``MPI_Sendrecv``: safe for ring shifts of any length (``send`` alone is a rendezvous beyond two slots)."""
        self._check(src, "sendrecv src"); self._check(dst, "sendrecv dst")
        L.check(self.lib.ewdml_coll_sendrecv(self._h, C.c_void_p(src.data_ptr()), _nbytes(src), to, C.c_void_p(dst.data_ptr()),
                                             _nbytes(dst), frm, tag, self._stream()), "coll_sendrecv")
        return dst

    @staticmethod
    def _f32(t: torch.Tensor) -> None:
        if t.dtype != torch.float32:
            raise TypeError("reductions are fp32 (cast first); got " + str(t.dtype))

    # ---- non-blocking forms ------------------------------------------------------------------------
    def _async(self, fn, *args, **kw) -> Request:
        if self._side is None:
            self._side = torch.cuda.Stream(device=self.device)
        self._side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self._side):
            res = fn(*args, **kw)
            ev = torch.cuda.Event()
            ev.record(self._side)
        self._last_async = ev
        return Request(ev, keep=(args, res))

    def iall_reduce(self, t, out=None, scale: float = 1.0) -> Request:
        return self._async(self.all_reduce, t, out, scale)

    def iall_gather(self, src, dst) -> Request:
        return self._async(self.all_gather, src, dst)

    def ibroadcast(self, t, root: int = 0) -> Request:
        return self._async(self.broadcast, t, root)

    def ireduce_scatter(self, src, dst, scale: float = 1.0) -> Request:
        return self._async(self.reduce_scatter, src, dst, scale)

    def iall_to_all(self, src, dst) -> Request:
        return self._async(self.all_to_all, src, dst)

    # ---- conveniences built on the above -----------------------------------------------------------
    def broadcast_parameters(self, tensors, root: int = 0) -> None:
        """``hvd.broadcast_parameters`` (``horvod_pytorch.py:187``): one flat message instead of one per tensor."""
        ts = [t for t in (tensors.values() if hasattr(tensors, "values") else tensors) if torch.is_tensor(t) and t.is_cuda]
        if not ts:
            return
        for dtype in sorted({t.dtype for t in ts}, key=str):
            grp = [t for t in ts if t.dtype == dtype]
            flat = torch.cat([t.detach().reshape(-1) for t in grp])
            self.broadcast(flat, root)
            off = 0
            for t in grp:
                t.detach().copy_(flat[off:off + t.numel()].view_as(t))
                off += t.numel()

    def metric_average(self, value: float) -> float:
        """``metric_average`` of the Horovod trainer (``horvod_pytorch.py:84-87``)."""
        t = torch.zeros(4, dtype=torch.float32, device=self.device)
        t[0] = float(value)
        self.all_reduce(t, scale=1.0 / self.world)
        return float(t[0].item())
