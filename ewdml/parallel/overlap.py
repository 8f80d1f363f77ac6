"""Bucketed, backward-overlapped gradient exchange (build step 6).

The reference's only attempt at overlapping communication with the backward pass is the dead
``LeNetSplit`` model: a hand-unrolled backward that ``Isend``s each layer's gradient as soon as it exists
(``src/model_ops/lenet.py:111-186``, mpi4py, import commented out).  Horovod gets the same effect with
autograd hooks and a 32 MB fusion buffer (``Horovod all reduce.ipynb`` cell 10).  This module provides that
capability for every model:

* :func:`partition_buckets` groups the flat buffer's segments (parameter tensors, forward order) into
  contiguous buckets of at least ``bucket_bytes``, built from the END of the model so the bucket that
  becomes ready first in backward is full-sized; many tiny BatchNorm vectors coalesce into their
  neighbours' buckets (ResNet-50 has 161 tensors).
* :class:`BucketScheduler` counts gradient arrivals per bucket and fires ``on_ready(bucket)`` the moment
  the last gradient of a bucket has been produced -- from ``register_post_accumulate_grad_hook`` on the
  torch compute path, from the native layer functions (which write gradients straight into the flat
  buffer and therefore never trigger autograd's hooks) through :func:`notify_grad_ready`.

What ``on_ready`` does is up to the exchange: the native allgather launches K1 (+K2) for that bucket on a
side CUDA stream (:class:`ewdml.parallel.native_exchange.NativeAllgather`), the gloo path starts an
asynchronous ``all_gather`` (:class:`ewdml.parallel.host_exchange.AllgatherExchange`).
"""
from __future__ import annotations

from typing import Callable, Dict, List, Sequence, Tuple

import torch

_ACTIVE: List["BucketScheduler"] = []


def partition_buckets(segments: Sequence, bucket_bytes: int) -> List[Tuple[int, int]]:
    """Half-open segment-index ranges ``[lo, hi)`` in forward order.  ``bucket_bytes <= 0``: one bucket."""
    n = len(segments)
    if bucket_bytes <= 0 or n <= 1:
        return [(0, n)]
    out, hi, acc = [], n, 0
    for i in range(n - 1, -1, -1):
        acc += segments[i].numel * 4
        if acc >= bucket_bytes:
            out.append((i, hi))
            hi, acc = i, 0
    if hi > 0:
        if out and acc < bucket_bytes // 2:          # a small remainder at the front joins the first layers' bucket
            lo, h = out.pop()
            out.append((0, h))
        else:
            out.append((0, hi))
    return out[::-1]


def notify_grad_ready(param: torch.nn.Parameter) -> None:
    """Called by the native layer functions after they wrote ``param``'s gradient into the flat buffer."""
    for s in _ACTIVE:
        s.param_ready(param)


class BucketScheduler:
    def __init__(self, params: Sequence[torch.nn.Parameter], buckets: Sequence[Tuple[int, int]],
                 on_ready: Callable[[int], None]):
        """``params[i]`` is the parameter of segment ``i``."""
        self.params, self.buckets, self.on_ready = list(params), list(buckets), on_ready
        self._bucket_of: Dict[int, int] = {}
        for b, (lo, hi) in enumerate(self.buckets):
            for i in range(lo, hi):
                self._bucket_of[id(self.params[i])] = b
        self._need = [hi - lo for lo, hi in self.buckets]
        self._left = list(self._need)
        self._seen: set = set()
        self._fired = [False] * len(self.buckets)
        self._handles = []
        self.active = False

    # -- wiring ---------------------------------------------------------------------------------------
    def install(self) -> "BucketScheduler":
        """Autograd hooks (torch compute path) + registration for the native layers' notifications."""
        for p in self.params:
            if p.requires_grad and hasattr(p, "register_post_accumulate_grad_hook"):
                self._handles.append(p.register_post_accumulate_grad_hook(self.param_ready))
        _ACTIVE.append(self)
        return self

    def remove(self) -> None:
        for h in self._handles:
            h.remove()
        self._handles.clear()
        if self in _ACTIVE:
            _ACTIVE.remove(self)

    # -- per step --------------------------------------------------------------------------------------
    def begin_step(self) -> None:
        self._left = list(self._need)
        self._seen.clear()
        self._fired = [False] * len(self.buckets)
        self.active = True

    def param_ready(self, p: torch.nn.Parameter) -> None:
        if not self.active:
            return
        k = id(p)
        b = self._bucket_of.get(k)
        if b is None or k in self._seen:
            return
        self._seen.add(k)
        self._left[b] -= 1
        if self._left[b] == 0:
            self._fire(b)

    def _fire(self, b: int) -> None:
        if not self._fired[b]:
            self._fired[b] = True
            self.on_ready(b)

    def flush(self) -> None:
        """After backward: fire whatever has not fired (parameters that received no gradient), last bucket first."""
        for b in range(len(self.buckets) - 1, -1, -1):
            self._fire(b)
        self.active = False
