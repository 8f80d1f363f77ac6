"""Gradient compressors: Horovod-style ``compress``/``decompress`` interface + oracles.

See :mod:`ewdml.compress.base` for the two calling conventions kept from the reference.
"""
from .base import Compressor, NoneCompressor, FP16Compressor
from .qsgd import QSGDCompressor, HorovodQSGDCompressor, qsgd_quantize, qsgd_dequantize, code_width
from .topk import TopKCompressor, sparsify, desparsify, topk_count
from .topk_qsgd import TopKQSGDCompressor


class Compression:
    """Namespace mirroring ``hvd.Compression`` (``horvod_pytorch.py:190-193``)."""
    none = NoneCompressor
    fp16 = FP16Compressor
    qsgd = HorovodQSGDCompressor


def build_compressor(name: str, qsgd_levels: int = 128, topk_ratio: float = 0.5, seed: int = 0):
    if name == "none":
        return NoneCompressor()
    if name == "qsgd":
        return QSGDCompressor(qsgd_levels, seed=seed)
    if name == "topk":
        return TopKCompressor(topk_ratio)
    if name == "topk-qsgd":
        return TopKQSGDCompressor(topk_ratio, qsgd_levels, seed=seed)
    raise ValueError(name)
