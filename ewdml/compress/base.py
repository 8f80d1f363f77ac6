"""Compressor interface.

Two calling conventions exist in the reference and both are kept:

* parameter-server flavour (``src/Compresssor/qsgd.py:12-40``): ``compress(t) -> payload`` where the
  payload is self-describing, ``decompress(payload) -> tensor``;
* Horovod flavour (``horovod_compression.py:11-43``, ``horovod.torch.compression.Compressor``):
  static ``compress(t) -> (tensor, ctx)``, ``decompress(tensor, ctx) -> tensor``.

:class:`Compressor` defines the Horovod-style pair plus ``payload_nbytes`` so the framework can
account exact wire bytes (the reference used ``sys.getsizeof(storage)``, which adds 72 B of Python
overhead per tensor -- ``distributed_worker.py:346``).
"""
from __future__ import annotations

from typing import Any, Tuple

import torch


class Compressor:
    """Horovod-style interface: ``compress(t) -> (payload, ctx)``; ``decompress(payload, ctx)``."""

    name = "base"

    def compress(self, tensor: torch.Tensor) -> Tuple[Any, Any]:
        raise NotImplementedError

    def decompress(self, payload: Any, ctx: Any) -> torch.Tensor:
        raise NotImplementedError

    def payload_nbytes(self, payload: Any, ctx: Any) -> int:
        """Bytes this payload occupies on the wire in the framework's packed format."""
        raise NotImplementedError

    # convenience: round trip
    def roundtrip(self, tensor: torch.Tensor) -> torch.Tensor:
        payload, ctx = self.compress(tensor)
        return self.decompress(payload, ctx).view_as(tensor)


class NoneCompressor(Compressor):
    """Identity (``hvd.Compression.none``; methods 1 and 3 of the report)."""

    name = "none"

    def compress(self, tensor):
        return tensor, None

    def decompress(self, payload, ctx=None):
        return payload

    def payload_nbytes(self, payload, ctx=None) -> int:
        return payload.numel() * payload.element_size()


class FP16Compressor(Compressor):
    """``hvd.Compression.fp16`` (mentioned at ``horvod_pytorch.py:190-191``)."""

    name = "fp16"

    def compress(self, tensor):
        if tensor.dtype.is_floating_point and tensor.dtype != torch.float16:
            return tensor.to(torch.float16), tensor.dtype
        return tensor, tensor.dtype

    def decompress(self, payload, ctx):
        return payload.to(ctx) if payload.dtype != ctx else payload

    def payload_nbytes(self, payload, ctx=None) -> int:
        return payload.numel() * payload.element_size()
