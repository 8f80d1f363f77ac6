"""Horovod-style front end: ``init / rank / size / allreduce / broadcast_parameters /
DistributedOptimizer(compression=...)`` (what ``horvod_pytorch.py:125-201`` uses from ``horovod.torch``).

Backed by ``torch.distributed`` (gloo on CPU, NCCL on GPU) for the generic tensor calls and by the
framework's compressors for gradients; ``init("native")`` instead brings up the peer-memory runtime and
routes ``allreduce`` / ``broadcast_parameters`` of CUDA tensors through the collective engine
(:mod:`ewdml.parallel.collectives`), with no ``torch.distributed`` process group at all.  Behavioural fix relative to the reference's Horovod path
(SURVEY.md 3.4): compressed gradients are all-gathered as (levels, norm) pairs and every pair is
dequantised with ITS OWN norm before averaging; the reference averages the levels of different
ranks and rescales by the local norm.
"""
from __future__ import annotations

import os
from typing import Iterable, Optional

import torch
import torch.distributed as dist

from ..compress import Compression, NoneCompressor

Average, Sum, Adasum = "average", "sum", "adasum"
_state = {"init": False}


def init(backend: Optional[str] = None) -> None:
    if backend == "native":
        if _state.get("engine") is None:
            from .collectives import CollectiveEngine
            from .peer import PeerGroup
            device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", os.environ.get("RANK", "0"))))
            torch.cuda.set_device(device)
            pg = PeerGroup.from_env(device)
            _state.update(pg=pg, engine=CollectiveEngine.create(pg), init=True)
        return
    if dist.is_initialized():
        _state["init"] = True
        return
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    dist.init_process_group(backend or "gloo", rank=rank, world_size=world)
    _state["init"] = True


def engine():
    """The native collective engine when ``init("native")`` was used, else None."""
    return _state.get("engine")


def shutdown() -> None:
    if _state.get("engine") is not None:
        _state["engine"].close()
        _state["pg"].close()
        _state.update(engine=None, pg=None)
    if dist.is_initialized():
        dist.destroy_process_group()


def rank() -> int:
    if _state.get("engine") is not None:
        return _state["engine"].rank
    return dist.get_rank() if dist.is_initialized() else 0


def size() -> int:
    if _state.get("engine") is not None:
        return _state["engine"].world
    return dist.get_world_size() if dist.is_initialized() else 1


def local_rank() -> int:
    return int(os.environ.get("LOCAL_RANK", rank()))


def local_size() -> int:
    return int(os.environ.get("LOCAL_WORLD_SIZE", size()))


def nccl_built() -> bool:
    return dist.is_nccl_available()


def _adasum_pair(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    dot, na, nb = (a * b).sum(), (a * a).sum(), (b * b).sum()
    ca = 1.0 - dot / (2 * na) if na > 0 else 1.0
    cb = 1.0 - dot / (2 * nb) if nb > 0 else 1.0
    return ca * a + cb * b


def allreduce(tensor: torch.Tensor, name: Optional[str] = None, op: str = Average,
              prescale_factor: float = 1.0, postscale_factor: float = 1.0) -> torch.Tensor:
    t = tensor.detach().clone() * prescale_factor
    eng = _state.get("engine")
    if eng is not None and t.is_cuda and op != Adasum:
        t32 = t.float().contiguous()
        eng.all_reduce(t32, scale=(1.0 / size() if op == Average else 1.0) * postscale_factor)
        return t32.to(tensor.dtype)
    if size() > 1:
        if op == Adasum:
            parts = [torch.empty_like(t) for _ in range(size())]
            dist.all_gather(parts, t)
            while len(parts) > 1:       # pairwise tree, as Horovod's recursive Adasum
                parts = [_adasum_pair(parts[i], parts[i + 1]) if i + 1 < len(parts) else parts[i]
                         for i in range(0, len(parts), 2)]
            t = parts[0]
        else:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            if op == Average:
                t /= size()
    return t * postscale_factor


def broadcast_parameters(params, root_rank: int = 0) -> None:
    items = params.items() if isinstance(params, dict) else params
    if size() == 1:
        return
    eng = _state.get("engine")
    if eng is not None:
        eng.broadcast_parameters([p for _, p in items if torch.is_tensor(p)], root_rank)
        return
    for _, p in items:
        if torch.is_tensor(p):
            dist.broadcast(p.data if hasattr(p, "data") else p, src=root_rank)


def broadcast_optimizer_state(optimizer: torch.optim.Optimizer, root_rank: int = 0) -> None:
    if size() == 1:
        return
    if _state.get("engine") is not None:
        # native runtime: no torch.distributed group exists; the (small, host-side) state travels through the rendezvous store
        import io
        pg = _state["pg"]
        _state["bcast_n"] = _state.get("bcast_n", 0) + 1
        key = f"hvd/optstate/{_state['bcast_n']}"
        if rank() == root_rank:
            buf = io.BytesIO()
            torch.save(optimizer.state_dict(), buf)
            pg.store.set(key, buf.getvalue())
        blob = pg.store.get(key, cap=1 << 28)
        if rank() != root_rank:
            optimizer.load_state_dict(torch.load(io.BytesIO(blob), map_location="cpu", weights_only=False))
        pg.host_barrier(key + "/done")
        return
    obj = [optimizer.state_dict() if rank() == root_rank else None]
    dist.broadcast_object_list(obj, src=root_rank)
    if rank() != root_rank:
        optimizer.load_state_dict(obj[0])


class _DistributedOptimizer:
    def __init__(self, optimizer, named_parameters=None, compression=Compression.none, op: str = Average,
                 gradient_predivide_factor: float = 1.0):
        self._opt = optimizer
        self._compression = compression() if isinstance(compression, type) else compression
        self._op = op
        self._predivide = gradient_predivide_factor
        self._params = [p for g in optimizer.param_groups for p in g["params"]]
        self._synchronized = False
        self.bytes_sent = 0

    def __getattr__(self, item):
        return getattr(self._opt, item)

    def zero_grad(self, *a, **k):
        return self._opt.zero_grad(*a, **k)

    def synchronize(self) -> None:
        W = size()
        for p in self._params:
            if p.grad is None:
                continue
            g = p.grad / self._predivide if self._predivide != 1.0 else p.grad
            if isinstance(self._compression, NoneCompressor) or W == 1:
                p.grad.copy_(allreduce(g, op=self._op, postscale_factor=self._predivide))
                continue
            payload, ctx = self._compression.compress(g)
            ctx_t = ctx if torch.is_tensor(ctx) else torch.zeros(1)
            pl = [torch.empty_like(payload) for _ in range(W)]
            cl = [torch.empty_like(ctx_t) for _ in range(W)]
            dist.all_gather(pl, payload.contiguous())
            dist.all_gather(cl, ctx_t.contiguous())
            self.bytes_sent += (W - 1) * (payload.numel() * payload.element_size() + 4)
            dec = [self._compression.decompress(q, c if torch.is_tensor(ctx) else ctx).view_as(p.grad) for q, c in zip(pl, cl)]
            if self._op == Adasum:
                while len(dec) > 1:
                    dec = [_adasum_pair(dec[i], dec[i + 1]) if i + 1 < len(dec) else dec[i] for i in range(0, len(dec), 2)]
                out = dec[0]
            else:
                out = sum(dec)
                if self._op == Average:
                    out = out / W
            p.grad.copy_(out * self._predivide)
        self._synchronized = True

    def step(self, closure=None):
        if not self._synchronized:
            self.synchronize()
        self._synchronized = False
        return self._opt.step(closure) if closure is not None else self._opt.step()


def DistributedOptimizer(optimizer, named_parameters=None, compression=Compression.none, op: str = Average,
                         gradient_predivide_factor: float = 1.0):
    return _DistributedOptimizer(optimizer, named_parameters, compression, op, gradient_predivide_factor)
