"""The allocation-free BatchNorm-statistics entry point, called through the C ABI.

``ewdml_bn_stats_static`` exists for users of libewdml.so that cannot allocate device memory on the
step path -- a C/C++ host loop with no torch caching allocator behind it, or a process in which the
GPU's memory is already reserved by something else, where a per-step cudaMalloc would stall the
device or fail.  Its scratch (2*C partial sums) is a fixed array inside the library, shared by every
call and not cleared between them by anyone but the call itself.  So the test runs it twice in a
row on different data (a stale-scratch bug would show on the second call), compares each result
with the ordinary entry points (``ewdml_colstats`` + ``ewdml_bn_finalize`` on a caller-provided
buffer), and checks the size limit.  This file sorts last on purpose: with ``-x`` every other test
has already run when it starts, and it also passes alone in a fresh process.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup():
    from ewdml.runtime import lib as L
    lib = L.load()
    vp, i32, i64, f32 = C.c_void_p, C.c_int, C.c_int64, C.c_float
    lib.ewdml_bn_stats_static.restype = i32
    lib.ewdml_bn_stats_static.argtypes = [vp, vp, vp, vp, vp, vp, i32, i64, f32, f32, vp]
    lib.ewdml_colstats.restype = i32
    lib.ewdml_colstats.argtypes = [vp, vp, i64, i32, i32, vp]
    lib.ewdml_bn_finalize.restype = i32
    lib.ewdml_bn_finalize.argtypes = [vp, vp, vp, vp, vp, vp, i32, i64, f32, f32, vp]
    return L, lib


def _p(t):
    return C.c_void_p(t.data_ptr())


def test_static_scratch_entry_matches_regular_path_twice_in_a_row():
    L, lib = _setup()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for call, (M, Cn, shift) in enumerate([(8192, 256, 3.0), (5000, 64, -1.5)]):
        g = torch.Generator(device="cpu").manual_seed(10 + call)
        y = (torch.randn(M, Cn, generator=g) * 2 + shift).to(dev).to(torch.bfloat16)
        gamma = (torch.rand(Cn, generator=g) + 0.5).to(dev)
        beta = torch.randn(Cn, generator=g).to(dev)
        rm_a, rv_a = torch.zeros(Cn, device=dev), torch.ones(Cn, device=dev)
        rm_b, rv_b = rm_a.clone(), rv_a.clone()
        stats_a = torch.empty(4 * Cn, device=dev)
        stats_b = torch.empty(4 * Cn, device=dev)
        # existing entry points, scratch supplied by the caller
        sums = torch.zeros(2 * Cn, device=dev)
        L.check(lib.ewdml_colstats(_p(y), _p(sums), M, Cn, 1, st))
        L.check(lib.ewdml_bn_finalize(_p(sums), _p(gamma), _p(beta), _p(rm_a), _p(rv_a), _p(stats_a), Cn, M, 0.1, 1e-5, st))
        # allocation-free entry point, scratch inside the library (still holding the previous call's sums)
        L.check(lib.ewdml_bn_stats_static(_p(y), _p(gamma), _p(beta), _p(rm_b), _p(rv_b), _p(stats_b), Cn, M, 0.1, 1e-5, st))
        torch.cuda.synchronize()
        assert torch.allclose(stats_b, stats_a, rtol=1e-4, atol=1e-5), f"call {call}"
        assert torch.allclose(rm_b, rm_a, rtol=1e-4, atol=1e-6) and torch.allclose(rv_b, rv_a, rtol=1e-4, atol=1e-6)
        # and both agree with a plain fp32 PyTorch reference of the same op
        yf = y.float()
        mean, var = yf.mean(0), yf.var(0, unbiased=False)
        assert torch.allclose(stats_b[:Cn], mean, rtol=1e-3, atol=1e-3)
        assert torch.allclose(stats_b[Cn:2 * Cn], torch.rsqrt(var + 1e-5), rtol=1e-3, atol=1e-4)


def test_static_scratch_entry_rejects_sizes_beyond_its_array():
    L, lib = _setup()
    dev = torch.device("cuda", 0)
    y = torch.zeros(4, 8192, device=dev, dtype=torch.bfloat16)
    v = torch.zeros(8192, device=dev)
    out = torch.zeros(4 * 8192, device=dev)
    rc = lib.ewdml_bn_stats_static(_p(y), _p(v), _p(v), None, None, _p(out), 8192, 4, 0.1, 1e-5,
                                   C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == -1                                                 # EWDML_ERR_ARG
    assert b"4096" in lib.ewdml_last_error()
