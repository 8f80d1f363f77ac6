"""ctypes binding of ``ewdml/_C/libewdml.so`` (declarations in ``include/ewdml.h``).

The library is built in-tree by :mod:`ewdml.build`.  On a machine with a GPU a missing library is
a hard error (no silent PyTorch fallback); on CPU-only machines :func:`available` is False and the
gloo path is used.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

MAX_PEERS = 16
TILE = 8192
MODE_DENSE, MODE_IDX16, MODE_BITMAP = 0, 1, 2
SEL_ALL, SEL_TOPK = 0, 1
VAL_QSGD, VAL_RAW32 = 0, 1
OUT_SGD, OUT_GRAD, OUT_ASSIGN = 0, 1, 2
AR_AUTO, AR_ONESHOT, AR_TWOSHOT, AR_NVLS = 0, 1, 2, 3
(COLL_BCAST, COLL_GATHER, COLL_SCATTER, COLL_ALLGATHER, COLL_ALLTOALL, COLL_REDUCE, COLL_REDUCE_SCATTER,
 COLL_ALLREDUCE, COLL_SEND, COLL_RECV, COLL_BARRIER) = range(11)
ALGO_AUTO, ALGO_ONESHOT, ALGO_TWOSHOT = 0, 1, 2
PHASE_PUT, PHASE_WAIT, PHASE_BOTH = 1, 2, 3
(CTR_LAUNCHES, CTR_SEND_CALLS, CTR_RECV_CALLS, CTR_ALLREDUCE_CALLS, CTR_BYTES_SENT, CTR_BYTES_RECV,
 CTR_COMPUTE_LAUNCHES) = range(7)

_LIB_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_C", "libewdml.so")
_lib: Optional[C.CDLL] = None


class Codec(C.Structure):
    _fields_ = [("mode", C.c_int), ("sel", C.c_int), ("val", C.c_int), ("levels", C.c_int),
                ("code_bits", C.c_int), ("topk_ratio", C.c_double)]


class Dests(C.Structure):
    _fields_ = [("n", C.c_int), ("slot", C.c_void_p * MAX_PEERS), ("flag", C.c_void_p * MAX_PEERS),
                ("mc_slot", C.c_void_p)]


class Sources(C.Structure):
    _fields_ = [("n", C.c_int), ("slot", C.c_void_p * MAX_PEERS), ("flag", C.c_void_p * MAX_PEERS)]


class Sgd(C.Structure):
    _fields_ = [("lr", C.c_float), ("momentum", C.c_float), ("dampening", C.c_float),
                ("weight_decay", C.c_float), ("nesterov", C.c_int), ("first_step", C.c_int)]


class Comm(C.Structure):
    _fields_ = [("rank", C.c_int), ("world", C.c_int), ("bufs", C.c_void_p * MAX_PEERS),
                ("peer_flags", C.c_void_p * MAX_PEERS), ("flags", C.c_void_p), ("mc_buf", C.c_void_p)]


class NativeError(RuntimeError):
    pass


def lib_path() -> str:
    return _LIB_PATH


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise NativeError(f"{_LIB_PATH} not found: run `python -m ewdml.build` (or __graft_entry__.build())")
    import torch  # noqa: F401  (initialises the CUDA libraries / primary context conventions first)
    lib = C.CDLL(_LIB_PATH, mode=C.RTLD_LOCAL)
    vp, i32, i64, u32, u64, f32 = C.c_void_p, C.c_int, C.c_int64, C.c_uint32, C.c_uint64, C.c_float
    P = C.POINTER
    sig = {
        "ewdml_last_error": (C.c_char_p, []),
        "ewdml_version": (i32, []),
        "ewdml_plan_create": (i32, [P(vp), i32, P(i64), P(i64), P(Codec)]),
        "ewdml_plan_destroy": (i32, [vp]),
        "ewdml_plan_slot_bytes": (i64, [vp]),
        "ewdml_plan_num_tiles": (i64, [vp]),
        "ewdml_plan_scratch_bytes": (i64, [vp]),
        "ewdml_plan_slot_layout": (i32, [vp, P(i64)]),
        "ewdml_send": (i32, [vp, vp, P(Dests), u32, u64, u32, u32, vp]),
        "ewdml_recv_apply": (i32, [vp, P(Sources), u32, f32, i32, P(Sgd), vp, vp, vp, vp, vp]),
        "ewdml_sgd_apply": (i32, [vp, vp, vp, vp, i64, P(Sgd), f32, vp]),
        "ewdml_state_advance": (i32, [vp, i32, vp]),
        "ewdml_send_state": (i32, [vp, vp, P(Dests), P(Dests), vp, u64, u32, vp]),
        "ewdml_recv_apply_state": (i32, [vp, P(Sources), P(Sources), vp, f32, i32, P(Sgd), vp, vp, vp, vp, vp]),
        "ewdml_sgd_apply_state": (i32, [vp, vp, vp, vp, i64, P(Sgd), f32, vp, vp]),
        "ewdml_allreduce_state": (i32, [P(Comm), i64, f32, i32, u32, vp, vp, vp]),
        "ewdml_plan_read_stats": (i32, [vp, P(i64), vp]),
        "ewdml_plan_read_seginfo": (i32, [vp, P(u32), vp]),
        "ewdml_plan_fallbacks": (i32, [vp, P(i64), vp]),
        "ewdml_allreduce": (i32, [P(Comm), i64, f32, i32, u32, vp, vp]),
        "ewdml_barrier": (i32, [P(Comm), u32, vp]),
        "ewdml_adasum_pair": (i32, [vp, vp, vp, i32, vp, vp, i64, vp, vp]),
        "ewdml_get_device_error": (i32, [P(u32), vp]),
        "ewdml_set_spin_limit": (i32, [u64]),
        "ewdml_heap_alloc": (i32, [P(vp), C.c_size_t]),
        "ewdml_heap_free": (i32, [vp]),
        "ewdml_heap_export": (i32, [vp, C.c_char_p]),
        "ewdml_heap_import": (i32, [C.c_char_p, P(vp)]),
        "ewdml_heap_unmap": (i32, [vp]),
        "ewdml_can_access_peer": (i32, [i32, i32, P(i32)]),
        "ewdml_sm_count": (i32, []),
        "ewdml_vmm_query": (i32, [P(i32)]),
        "ewdml_vmm_granularity": (i32, [i32, P(C.c_size_t)]),
        "ewdml_vmm_alloc": (i32, [P(vp), C.c_size_t, P(vp), P(i32)]),
        "ewdml_vmm_import": (i32, [P(vp), i32, C.c_size_t, P(vp)]),
        "ewdml_vmm_close_fd": (i32, [i32]),
        "ewdml_vmm_free": (i32, [vp]),
        "ewdml_mc_create": (i32, [P(vp), i32, C.c_size_t, P(i32)]),
        "ewdml_mc_import": (i32, [P(vp), i32, C.c_size_t]),
        "ewdml_mc_add_device": (i32, [vp]),
        "ewdml_mc_bind_map": (i32, [vp, vp, P(vp)]),
        "ewdml_fdx_open": (i32, [P(vp), C.c_char_p, i32, i32]),
        "ewdml_fdx_send": (i32, [vp, i32, i32, u64, C.c_double]),
        "ewdml_fdx_recv": (i32, [vp, u64, i32, P(i32), P(i32), C.c_double]),
        "ewdml_fdx_close": (i32, [vp]),
        "ewdml_store_create": (i32, [P(vp), C.c_char_p, i32, i32, i32, C.c_double]),
        "ewdml_store_set": (i32, [vp, C.c_char_p, C.c_char_p, C.c_size_t]),
        "ewdml_store_get": (i32, [vp, C.c_char_p, C.c_char_p, C.c_size_t, P(C.c_size_t)]),
        "ewdml_store_barrier": (i32, [vp, C.c_char_p]),
        "ewdml_store_destroy": (i32, [vp]),
        "ewdml_launch": (i32, [i32, C.c_char_p, i32, P(C.c_char_p), C.c_double, P(i32)]),
        "ewdml_counter_get": (i64, [i32]),
        "ewdml_counter_add": (None, [i32, i64]),
        "ewdml_counters_reset": (None, []),
        "ewdml_peer_matrix_add": (None, [i32, i64]),
        "ewdml_peer_matrix_get": (i32, [P(i64)]),
        "ewdml_set_trace": (None, [i32]),
        "ewdml_symheap_create": (i32, [P(vp), i64, i64]),
        "ewdml_symheap_alloc": (i64, [vp, i64]),
        "ewdml_symheap_free": (i32, [vp, i64]),
        "ewdml_symheap_used": (i64, [vp]),
        "ewdml_symheap_peak": (i64, [vp]),
        "ewdml_symheap_largest_free": (i64, [vp]),
        "ewdml_symheap_destroy": (i32, [vp]),
        "ewdml_engine_window_bytes": (i64, [i32, i64]),
        "ewdml_engine_flag_bytes": (i64, []),
        "ewdml_engine_create": (i32, [P(vp), i32, i32, P(vp), P(vp), i64]),
        "ewdml_engine_destroy": (i32, [vp]),
        "ewdml_engine_set_algo": (i32, [vp, i32, i32]),
        "ewdml_engine_select_algo": (i32, [vp, i32, i64]),
        "ewdml_engine_calls": (i64, [vp, i32]),
        "ewdml_engine_slot_payload": (i64, [vp]),
        "ewdml_coll_op_name": (C.c_char_p, [i32]),
        "ewdml_coll": (i32, [vp, i32, vp, vp, i64, i32, f32, i32, vp]),
        "ewdml_coll_send": (i32, [vp, vp, i64, i32, i32, vp]),
        "ewdml_coll_recv": (i32, [vp, vp, i64, i32, i32, vp]),
        "ewdml_coll_sendrecv": (i32, [vp, vp, i64, i32, vp, i64, i32, i32, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().ewdml_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'ewdml call'} failed (code {rc}): {msg}")


def launches() -> int:
    """Kernels launched by this library in this process (bench.py's ``gpu_launches``)."""
    lib = load()
    return int(lib.ewdml_counter_get(CTR_LAUNCHES)) + int(lib.ewdml_counter_get(CTR_COMPUTE_LAUNCHES))


class Store:
    """TCP rendezvous store (``ewdml_store_*``)."""

    def __init__(self, addr: str, port: int, rank: int, world: int, timeout_s: float = 60.0):
        self._lib = load()
        self._h = C.c_void_p()
        check(self._lib.ewdml_store_create(C.byref(self._h), addr.encode(), port, rank, world, timeout_s),
              "store_create")
        self.rank, self.world = rank, world

    def set(self, key: str, val: bytes) -> None:
        check(self._lib.ewdml_store_set(self._h, key.encode(), val, len(val)), "store_set")

    def get(self, key: str, cap: int = 1 << 16) -> bytes:
        """Blocking get.  ``cap`` is only the first guess: a larger value is fetched again into a buffer of its real size."""
        buf = C.create_string_buffer(cap)
        n = C.c_size_t()
        rc = self._lib.ewdml_store_get(self._h, key.encode(), buf, cap, C.byref(n))
        if rc == -4 and n.value > cap:                 # EWDML_ERR_OVERFLOW: the length is known now
            cap = n.value
            buf = C.create_string_buffer(cap)
            rc = self._lib.ewdml_store_get(self._h, key.encode(), buf, cap, C.byref(n))
        check(rc, "store_get")
        return buf.raw[:n.value]

    def barrier(self, name: str = "default") -> None:
        check(self._lib.ewdml_store_barrier(self._h, name.encode()), "store_barrier")

    def all_gather(self, name: str, val: bytes):
        self.set(f"{name}/{self.rank}", val)
        return [self.get(f"{name}/{r}") for r in range(self.world)]

    def close(self) -> None:
        if self._h:
            self._lib.ewdml_store_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def launch(nproc: int, argv, master_addr: str = "127.0.0.1", master_port: int = 29500, grace_s: float = 5.0):
    """Run ``argv`` as ``nproc`` local ranks under the native watchdog; returns (status, failed_rank)."""
    lib = load()
    arr = (C.c_char_p * (len(argv) + 1))(*[a.encode() for a in argv], None)
    failed = C.c_int(-1)
    rc = lib.ewdml_launch(nproc, master_addr.encode(), master_port, arr, grace_s, C.byref(failed))
    return rc, failed.value
