"""Process group over the symmetric peer heap (the ``ProcessGroup`` object of SURVEY.md 2E).

A :class:`PeerGroup` knows (rank, world, device), performs the start-up rendezvous and hands out
*symmetric buffers*: every rank allocates the same number of bytes, IPC handles are exchanged once
through the native TCP store (or ``torch.distributed`` when it is already initialised) and every
peer's allocation is mapped into this process, after which kernels address peer memory with plain
pointers over NVLink.  This replaces communicator construction + PMIx modex + BTL endpoint set-up
(``ompi:ompi/runtime/ompi_mpi_init.c:428-800``) and the smcuda IPC handshake.

:class:`LocalPeerGroup` emulates ``world`` ranks inside ONE process on ONE GPU (every "peer" buffer
is a local allocation).  It exists for single-GPU tests: kernels that wait on each other must never
run as separate launches on one GPU (B200_PROFILING.md), so the emulation runs all senders first
and all receivers afterwards on one stream.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

import torch

from ..runtime import lib as L


class _CudaArray:
    """Minimal ``__cuda_array_interface__`` carrier so torch can wrap raw device memory."""

    def __init__(self, ptr: int, nbytes: int, owner=None):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False),
                                         "version": 2, "strides": None}
        self._owner = owner


def wrap_device_memory(ptr: int, nbytes: int, device: torch.device, owner=None) -> torch.Tensor:
    with torch.cuda.device(device):
        t = torch.as_tensor(_CudaArray(ptr, nbytes, owner), device=device)
    return t


class SymmetricBuffer:
    """``ptrs[r]`` = address of rank r's copy as mapped in this process; ``local`` = own bytes tensor."""

    def __init__(self, name: str, nbytes: int, ptrs: List[int], rank: int, device: torch.device, owner, mc_ptr: int = 0):
        self.name, self.nbytes, self.ptrs, self.rank = name, nbytes, ptrs, rank
        self.local = wrap_device_memory(ptrs[rank], nbytes, device, owner)
        # NVLS: address of the multicast mapping of this buffer (a store to mc_ptr + x lands at offset x of EVERY rank's
        # copy, a multimem.ld_reduce from it returns the sum over all copies); 0 when the heap has none
        self.mc_ptr = mc_ptr

    def view(self, dtype: torch.dtype, offset_bytes: int = 0, numel: Optional[int] = None) -> torch.Tensor:
        es = torch.empty((), dtype=dtype).element_size()
        end = self.nbytes if numel is None else offset_bytes + numel * es
        return self.local[offset_bytes:end].view(dtype)


class _Alloc:
    def __init__(self, lib, ptr):
        self.lib, self.ptr = lib, ptr

    def __del__(self):
        try:
            if self.ptr:
                self.lib.ewdml_heap_free(C.c_void_p(self.ptr))
        except Exception:
            pass


class _VmmAlloc:
    """Owner of one VMM-backed symmetric buffer: own allocation, imported peer mappings, multicast object."""

    def __init__(self, lib, own, imported, mc):
        self.lib, self.own, self.imported, self.mc = lib, own, imported, mc

    def release(self):
        for h in ([self.mc] if self.mc else []) + list(self.imported) + ([self.own] if self.own else []):
            try:
                self.lib.ewdml_vmm_free(h)
            except Exception:
                pass
        self.own, self.imported, self.mc = None, [], None

    def __del__(self):
        self.release()


class PeerGroup:
    """``mode``: ``"ipc"`` (cudaMalloc + CUDA IPC handles through the store) or ``"vmm"`` (cuMemCreate allocations whose
    POSIX file descriptors travel over a UNIX socket, which is what NVLS multicast objects need); ``multicast`` says
    whether buffers come with a multicast mapping (``SymmetricBuffer.mc_ptr``).  ``EWDML_HEAP=auto|vmm|ipc`` and
    ``EWDML_MULTICAST=0|1`` choose; auto falls back to IPC unless EVERY rank reports VMM + fd-handle support."""

    def __init__(self, rank: int, world: int, device: torch.device, store: Optional[L.Store] = None):
        self.lib = L.load()
        self.rank, self.world, self.device = rank, world, device
        self.store = store
        self._allocs = []
        self._n = 0
        self.mode, self.multicast, self._fdx = "ipc", False, None
        torch.cuda.set_device(device)
        if world > L.MAX_PEERS:
            raise ValueError(f"world size {world} exceeds EWDML_MAX_PEERS={L.MAX_PEERS}")
        if world > 1:
            self._negotiate_transport()

    # -- transport negotiation: every rank reports what its device/driver can do; the group takes the common subset ----
    def _negotiate_transport(self) -> None:
        from ..config import registry
        want = str(registry.get("heap")).lower()
        if want == "ipc":
            return
        q = (C.c_int * 3)()
        self.lib.ewdml_vmm_query(q)
        mine = bytes([int(bool(q[0] and q[2])), int(bool(q[1]) and bool(registry.get("multicast")))])
        import uuid
        job = uuid.uuid4().hex[:16].encode()
        every = self._all_gather_bytes("transport", mine + job)
        if not all(e[0] for e in every):
            if want == "vmm":
                raise L.NativeError("EWDML_HEAP=vmm but a rank lacks VMM / POSIX-fd handle support")
            return
        fdx = C.c_void_p()
        rc = self.lib.ewdml_fdx_open(C.byref(fdx), every[0][2:], self.rank, self.world)
        ok = self._all_gather_bytes("fdx_open", bytes([int(rc == 0)]))
        if not all(o[0] for o in ok):
            if rc == 0:
                self.lib.ewdml_fdx_close(fdx)
            if want == "vmm":
                raise L.NativeError("EWDML_HEAP=vmm but the fd channel could not be opened on every rank")
            return
        self._fdx, self.mode = fdx, "vmm"
        self.multicast = all(e[1] for e in every)

    @classmethod
    def from_env(cls, device: torch.device) -> "PeerGroup":
        rank = int(os.environ.get("RANK", "0"))
        world = int(os.environ.get("WORLD_SIZE", "1"))
        store = None
        if world > 1:
            addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
            port = int(os.environ.get("EWDML_STORE_PORT", int(os.environ.get("MASTER_PORT", "29500")) + 17))
            if "EWDML_STORE_PORT" not in os.environ:
                # when torch.distributed is already up, let rank 0 pick a port that is free right now instead of
                # trusting MASTER_PORT + 17
                import torch.distributed as dist
                if dist.is_available() and dist.is_initialized():
                    obj = [None]
                    if rank == 0:
                        import socket
                        s = socket.socket()
                        s.bind(("", 0))
                        obj[0] = s.getsockname()[1]
                        s.close()
                    dist.broadcast_object_list(obj, src=0, device=device if dist.get_backend() == "nccl" else None)
                    port = int(obj[0])
            store = L.Store(addr, port, rank, world, timeout_s=120.0)
        return cls(rank, world, device, store)

    def _all_gather_bytes(self, name: str, val: bytes) -> List[bytes]:
        if self.world == 1:
            return [val]
        if self.store is not None:
            return self.store.all_gather(name, val)
        import torch.distributed as dist
        out = [None] * self.world
        dist.all_gather_object(out, val)
        return out

    def host_barrier(self, name: str = "pg") -> None:
        if self.world == 1:
            return
        if self.store is not None:
            self.store.barrier(name)
        else:
            import torch.distributed as dist
            dist.barrier()

    def _agree(self, key: str, ok: bool) -> bool:
        """True iff `ok` on every rank (a collective vote through the store)."""
        return all(v[0] for v in self._all_gather_bytes(key, bytes([int(bool(ok))])))

    def _alloc_vmm(self, name: str, nbytes: int) -> SymmetricBuffer:
        lib, me, W = self.lib, self.rank, self.world
        gran = C.c_size_t()
        L.check(lib.ewdml_vmm_granularity(W if self.multicast else 1, C.byref(gran)), "vmm_granularity")
        padded = (nbytes + gran.value - 1) // gran.value * gran.value
        own, base, fd = C.c_void_p(), C.c_void_p(), C.c_int(-1)
        L.check(lib.ewdml_vmm_alloc(C.byref(own), padded, C.byref(base), C.byref(fd)), f"vmm_alloc({name}, {padded})")
        self._n += 1
        n = self._n
        for r in range(W):
            if r != me:
                L.check(lib.ewdml_fdx_send(self._fdx, r, fd.value, (n << 8) | 0, 120.0), f"fdx_send(mem -> rank {r})")
        ptrs, imported = [], []
        for r in range(W):
            if r == me:
                ptrs.append(base.value)
                continue
            pfd, h, mapped = C.c_int(-1), C.c_void_p(), C.c_void_p()
            L.check(lib.ewdml_fdx_recv(self._fdx, (n << 8) | 0, r, C.byref(pfd), None, 120.0), f"fdx_recv(mem <- rank {r})")
            L.check(lib.ewdml_vmm_import(C.byref(h), pfd.value, padded, C.byref(mapped)), f"vmm_import(rank {r})")
            lib.ewdml_vmm_close_fd(pfd.value)
            ptrs.append(mapped.value)
            imported.append(h)
        self.host_barrier(f"vmm/{n}/mapped")
        lib.ewdml_vmm_close_fd(fd.value)
        mc_h, mc_ptr = None, 0
        if self.multicast:
            # rank 0 creates the multicast object; every phase is voted on so that a failure (no fabric manager, object
            # limit) turns multicast off on ALL ranks instead of leaving some waiting
            mc, mfd = C.c_void_p(), C.c_int(-1)
            rc = lib.ewdml_mc_create(C.byref(mc), W, padded, C.byref(mfd)) if me == 0 else 0
            err0 = lib.ewdml_last_error().decode(errors="replace") if rc else ""
            if self._agree(f"mc/{n}/create", rc == 0):
                if me == 0:
                    for r in range(1, W):
                        L.check(lib.ewdml_fdx_send(self._fdx, r, mfd.value, (n << 8) | 1, 120.0), "fdx_send(mc)")
                else:
                    L.check(lib.ewdml_fdx_recv(self._fdx, (n << 8) | 1, 0, C.byref(mfd), None, 120.0), "fdx_recv(mc)")
                    rc = lib.ewdml_mc_import(C.byref(mc), mfd.value, padded)
                rc = rc or lib.ewdml_mc_add_device(mc)
                if self._agree(f"mc/{n}/add", rc == 0):
                    mp = C.c_void_p()
                    rc = lib.ewdml_mc_bind_map(mc, own, C.byref(mp))
                    if self._agree(f"mc/{n}/bind", rc == 0):
                        mc_h, mc_ptr = mc, mp.value
                lib.ewdml_vmm_close_fd(mfd.value)
            if not mc_ptr:
                from ..utils.metrics import logger
                if me == 0:
                    logger.warning("NVLS multicast unavailable (%s); pushes stay unicast", err0 or lib.ewdml_last_error().decode(errors="replace"))
                if mc:
                    lib.ewdml_vmm_free(mc)
                self.multicast = False
        owner = _VmmAlloc(lib, own, imported, mc_h)
        buf = SymmetricBuffer(name, padded, ptrs, me, self.device, owner, mc_ptr)
        buf._vmm = owner
        self._allocs.append(buf)
        return buf

    def alloc(self, name: str, nbytes: int) -> SymmetricBuffer:
        nbytes = (nbytes + 255) // 256 * 256
        if self.mode == "vmm":
            return self._alloc_vmm(name, nbytes)
        base = C.c_void_p()
        L.check(self.lib.ewdml_heap_alloc(C.byref(base), nbytes), f"heap_alloc({name}, {nbytes})")
        owner = _Alloc(self.lib, base.value)
        handle = C.create_string_buffer(64)
        L.check(self.lib.ewdml_heap_export(base, handle), "heap_export")
        self._n += 1
        handles = self._all_gather_bytes(f"ipc/{self._n}/{name}", handle.raw)
        ptrs = []
        for r, h in enumerate(handles):
            if r == self.rank:
                ptrs.append(base.value)
            else:
                mapped = C.c_void_p()
                L.check(self.lib.ewdml_heap_import(h, C.byref(mapped)), f"heap_import(rank {r})")
                ptrs.append(mapped.value)
        self.host_barrier(f"alloc/{self._n}")
        buf = SymmetricBuffer(name, nbytes, ptrs, self.rank, self.device, owner)
        self._allocs.append(buf)
        return buf

    def close(self) -> None:
        for b in self._allocs:
            if getattr(b, "_vmm", None) is not None:
                continue            # VMM mappings live as long as the tensors that view them (freed by their owner object)
            for r, p in enumerate(b.ptrs):
                if r != self.rank and p:
                    self.lib.ewdml_heap_unmap(C.c_void_p(p))
        self._allocs.clear()
        if self._fdx is not None:
            self.lib.ewdml_fdx_close(self._fdx)
            self._fdx = None
        if self.store is not None:
            self.store.close()
            self.store = None


class SymmetricHeap:
    """One peer-mapped region + the native deterministic sub-allocator (``ewdml_symheap_*``): after the single IPC
    exchange that maps the region, ``alloc`` is a purely local call -- every rank performs the same sequence of
    allocations, so an object sits at the same offset in every rank's region (OSHMEM ``shmalloc`` semantics,
    ``oshmem/mca/memheap``)."""

    def __init__(self, pg, nbytes: int, name: str = "symheap", align: int = 256):
        self.pg, self.lib = pg, L.load()
        self.region = pg.alloc(name, nbytes)          # PeerGroup: one SymmetricBuffer; LocalPeerGroup: one per rank
        self._regions = self.region if isinstance(self.region, list) else [self.region]
        self._h = C.c_void_p()
        L.check(self.lib.ewdml_symheap_create(C.byref(self._h), self._regions[0].nbytes, align), "symheap_create")
        self._live = {}

    def alloc(self, name: str, nbytes: int):
        off = int(self.lib.ewdml_symheap_alloc(self._h, int(nbytes)))
        if off < 0:
            raise L.NativeError(self.lib.ewdml_last_error().decode(errors="replace"))
        bufs = []
        for reg in self._regions:
            b = SymmetricBuffer.__new__(SymmetricBuffer)
            b.name, b.nbytes, b.rank = name, (nbytes + 255) // 256 * 256, reg.rank
            b.ptrs = [p + off for p in reg.ptrs]
            b.mc_ptr = reg.mc_ptr + off if getattr(reg, "mc_ptr", 0) else 0
            b.local = reg.local[off:off + b.nbytes]
            b.heap_offset = off
            bufs.append(b)
        out = bufs if isinstance(self.region, list) else bufs[0]
        self._live[off] = out
        return out

    def free(self, buf) -> None:
        b = buf[0] if isinstance(buf, list) else buf
        L.check(self.lib.ewdml_symheap_free(self._h, b.heap_offset), "symheap_free")
        self._live.pop(b.heap_offset, None)

    @property
    def used(self) -> int:
        return int(self.lib.ewdml_symheap_used(self._h))

    @property
    def largest_free(self) -> int:
        return int(self.lib.ewdml_symheap_largest_free(self._h))

    def close(self) -> None:
        if self._h:
            self.lib.ewdml_symheap_destroy(self._h)
            self._h = C.c_void_p()


class LocalPeerGroup:
    """``world`` emulated ranks in one process on one GPU; ``alloc`` returns one buffer per rank."""

    def __init__(self, world: int, device: torch.device):
        self.lib = L.load()
        self.world, self.device = world, device
        torch.cuda.set_device(device)

    def alloc(self, name: str, nbytes: int) -> List[SymmetricBuffer]:
        nbytes = (nbytes + 255) // 256 * 256
        owners, ptrs = [], []
        for _ in range(self.world):
            base = C.c_void_p()
            L.check(self.lib.ewdml_heap_alloc(C.byref(base), nbytes), f"heap_alloc({name})")
            owners.append(_Alloc(self.lib, base.value))
            ptrs.append(base.value)
        return [SymmetricBuffer(name, nbytes, list(ptrs), r, self.device, owners[r]) for r in range(self.world)]
