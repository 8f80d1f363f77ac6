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
        _lib.ewdml_gemm_bf16_tn.restype = i32
        _lib.ewdml_gemm_bf16_tn.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp, f32, i32, vp]
    return _lib


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t: Optional[torch.Tensor]):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


OUT_BF16, OUT_F32, OUT_ACCUM = 0, 1, 2


def gemm_tn(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None, out_mode: int = OUT_BF16,
            bias: Optional[torch.Tensor] = None, alpha: float = 1.0, split_k: int = 0) -> torch.Tensor:
    """``out (op)= alpha * a @ b.T (+ bias)`` with a:[M,K], b:[N,K] bf16 (row pitch may exceed K)."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.dim() == 2 and b.dim() == 2
    assert a.stride(1) == 1 and b.stride(1) == 1 and a.shape[1] == b.shape[1]
    M, K = a.shape
    N = b.shape[0]
    if out is None:
        out = torch.empty(M, N, dtype=torch.bfloat16 if out_mode == OUT_BF16 else torch.float32, device=a.device)
        if out_mode == OUT_ACCUM:
            out.zero_()
    assert out.stride(1) == 1
    L.check(_l().ewdml_gemm_bf16_tn(_p(a), _p(b), _p(out), M, N, K, a.stride(0), b.stride(0), out.stride(0), out_mode,
                                    _p(bias), float(alpha), split_k, _stream()), "gemm_bf16_tn")
    return out
