"""Property tests (hypothesis) of the host-side invariants: bucket partitioning, the symmetric-heap sub-allocator, the byte
codec round trip (pack / unpack bit-exactness), exact-k selection and the QSGD level range."""
import ctypes as C
from collections import namedtuple

import pytest
import torch
from hypothesis import given, settings, strategies as st

from ewdml.compress.qsgd import qsgd_dequantize, qsgd_quantize
from ewdml.compress.topk import sparsify, topk_count
from ewdml.parallel.flat import Segment
from ewdml.parallel.host_codec import HostCodec, _pack_levels, _unpack_levels
from ewdml.parallel.overlap import partition_buckets

Seg = namedtuple("Seg", "numel")


@settings(max_examples=200, deadline=None)
@given(st.lists(st.integers(1, 5_000_000), min_size=1, max_size=200), st.integers(0, 64 << 20))
def test_partition_buckets_covers_contiguously(sizes, bucket_bytes):
    segs = [Seg(n) for n in sizes]
    b = partition_buckets(segs, bucket_bytes)
    assert b[0][0] == 0 and b[-1][1] == len(segs)
    assert all(lo < hi for lo, hi in b) and all(b[i][1] == b[i + 1][0] for i in range(len(b) - 1))
    if bucket_bytes > 0:
        for lo, hi in b[1:]:                                   # every bucket except the first reaches the target size
            assert sum(s.numel for s in segs[lo:hi]) * 4 >= bucket_bytes
    else:
        assert b == [(0, len(segs))]


@settings(max_examples=60, deadline=None)
@given(st.lists(st.tuples(st.booleans(), st.integers(1, 50_000)), min_size=1, max_size=120))
def test_symmetric_heap_never_overlaps_and_is_deterministic(ops):
    from ewdml.runtime import lib as L
    lib = L.load()

    def run():
        h = C.c_void_p()
        L.check(lib.ewdml_symheap_create(C.byref(h), 1 << 20, 256))
        live, trace = {}, []
        for do_alloc, n in ops:
            if do_alloc or not live:
                off = lib.ewdml_symheap_alloc(h, n)
                trace.append(off)
                if off >= 0:
                    size = (n + 255) // 256 * 256
                    assert off % 256 == 0 and off + size <= 1 << 20
                    assert all(off + size <= o or o + s <= off for o, s in live.items())      # no overlap
                    live[off] = size
            else:
                off = sorted(live)[n % len(live)]
                L.check(lib.ewdml_symheap_free(h, off))
                del live[off]
                trace.append(-2)
            assert lib.ewdml_symheap_used(h) == sum(live.values())
        lib.ewdml_symheap_destroy(h)
        return trace

    assert run() == run()


@settings(max_examples=100, deadline=None)
@given(st.integers(1, 3000), st.sampled_from([1, 2, 4, 16, 64, 128, 256, 1024]), st.integers(0, 2 ** 31 - 1))
def test_level_packing_round_trip_is_bit_exact(n, s, seed):
    g = torch.Generator().manual_seed(seed)
    q = torch.randint(-s, s + 1, (n,), generator=g).float()
    assert torch.equal(_unpack_levels(_pack_levels(q, s), n, s), q)


@settings(max_examples=60, deadline=None)
@given(st.integers(1, 5000), st.floats(0.001, 1.0), st.integers(0, 2 ** 31 - 1))
def test_topk_selects_exactly_k_largest(n, ratio, seed):
    g = torch.Generator().manual_seed(seed)
    t = torch.randn(n, generator=g)
    vals, idx = sparsify(t, ratio)
    k = topk_count(n, ratio)
    assert vals.numel() == k == max(1, int(n * ratio)) and idx.unique().numel() == k
    assert torch.equal(vals, t[idx])
    if k < n:
        rest = torch.ones(n, dtype=torch.bool)
        rest[idx] = False
        assert vals.abs().min() >= t[rest].abs().max()


@settings(max_examples=60, deadline=None)
@given(st.integers(1, 4000), st.sampled_from([1, 4, 64, 128, 255, 256]), st.integers(0, 2 ** 31 - 1), st.sampled_from([0.0, 1e-30, 1.0, 1e20]))
def test_qsgd_levels_in_range_and_error_bounded(n, s, seed, scale):
    g = torch.Generator().manual_seed(seed)
    t = torch.randn(n, generator=g) * scale
    u = torch.rand(n, generator=g)
    q, norm = qsgd_quantize(t, s, u)
    assert q.abs().max() <= s and torch.equal(q, q.round())
    d = qsgd_dequantize(q, norm, s)
    assert torch.isfinite(d).all()
    if float(norm) > 0 and torch.isfinite(norm):
        assert ((d - t).abs() <= norm / s * (1 + 1e-5) + 1e-30).all()          # one level of error at most
        assert torch.equal(torch.sign(d)[d != 0], torch.sign(t)[d != 0])
    else:
        assert float(d.abs().sum()) == 0.0                                       # zero-norm tensors encode to zeros


@settings(max_examples=40, deadline=None)
@given(st.lists(st.integers(1, 3000), min_size=1, max_size=6), st.sampled_from(["none", "qsgd", "topk", "topk-qsgd"]),
       st.floats(0.01, 1.0), st.integers(0, 1000))
def test_host_codec_message_size_and_decode_support(sizes, comp, ratio, seed):
    segs, off = [], 0
    for i, n in enumerate(sizes):
        segs.append(Segment(f"p{i}", off, n, torch.Size([n]), False))
        off = (off + n + 31) // 32 * 32
    g = torch.Generator().manual_seed(seed)
    flat = torch.randn(max(off, 32), generator=g)
    codec = HostCodec(segs, comp, 128, ratio, seed=seed, rank=1)
    msg = codec.encode(flat, step=seed)
    assert msg.numel() == codec.nbytes and msg.dtype == torch.uint8
    dense = codec.decode(msg, total=flat.numel())
    for sg in segs:
        d, t = dense[sg.offset:sg.offset + sg.numel], flat[sg.offset:sg.offset + sg.numel]
        if comp == "none":
            assert torch.equal(d, t)
        if comp in ("topk", "topk-qsgd"):
            assert int((d != 0).sum()) <= topk_count(sg.numel, ratio)
        if comp == "topk":
            assert torch.equal(d[d != 0], t[d != 0])
    assert torch.equal(codec.encode(flat, step=seed), msg)                       # counter-based RNG: reproducible
