"""Hand-written sm_100a compute ops (K5-K9): thin autograd wrappers over ``libewdml.so``.

Activations are bf16 in NHWC (torch ``channels_last``) order, weights are read from the bf16 shadow
copy that the fused optimizer kernel maintains next to the fp32 master parameters, and weight
gradients are accumulated by the kernels straight into the flat fp32 gradient buffer.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from ..runtime import lib as L

_lib = None


def _l():
    global _lib
    if _lib is None:
        _lib = L.load()
        vp, i32, f32 = C.c_void_p, C.c_int, C.c_float
        _lib.ewdml_gemm_bf16.restype = i32
        _lib.ewdml_gemm_bf16.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp, f32, i32, vp]
    return _lib


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t: Optional[torch.Tensor]):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


OUT_BF16, OUT_F32, OUT_ACCUM = 0, 1, 2


def gemm(a: torch.Tensor, b: torch.Tensor, a_mn: bool = False, b_mn: bool = False,
         out: Optional[torch.Tensor] = None, out_mode: int = OUT_BF16, bias: Optional[torch.Tensor] = None,
         alpha: float = 1.0, split_k: int = 0) -> torch.Tensor:
    """tcgen05 GEMM ``out (op)= alpha * op(a) @ op(b) (+ bias)`` on bf16 row-major matrices.

    ``a_mn=False``: a is [M, K];  ``a_mn=True``: a is [K, M] (used as its transpose).
    ``b_mn=False``: b is [N, K];  ``b_mn=True``: b is [K, N].
    Row pitches may exceed the logical width (views of wider matrices are fine).
    """
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.dim() == 2 and b.dim() == 2
    assert a.stride(1) == 1 and b.stride(1) == 1
    (K, M) = a.shape if a_mn else (a.shape[1], a.shape[0])
    (Kb, N) = b.shape if b_mn else (b.shape[1], b.shape[0])
    assert K == Kb, (a.shape, b.shape, a_mn, b_mn)
    if out is None:
        out = torch.empty(M, N, dtype=torch.bfloat16 if out_mode == OUT_BF16 else torch.float32, device=a.device)
        if out_mode == OUT_ACCUM:
            out.zero_()
    assert out.stride(-1) == 1
    ldc = out.stride(0) if out.dim() == 2 else N
    L.check(_l().ewdml_gemm_bf16(_p(a), _p(b), _p(out), M, N, K, a.stride(0), b.stride(0), ldc, int(a_mn), int(b_mn),
                                 out_mode, _p(bias), float(alpha), split_k, _stream()), "gemm_bf16")
    return out


def gemm_tn(a, b, **kw):
    return gemm(a, b, False, False, **kw)
