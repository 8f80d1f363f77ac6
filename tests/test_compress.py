"""Oracle tests for QSGD / top-k / top-k->QSGD (implication (a))."""
import math

import numpy as np
import pytest
import torch

from ewdml.compress import (QSGDCompressor, TopKCompressor, TopKQSGDCompressor, HorovodQSGDCompressor,
                            NoneCompressor, FP16Compressor, Compression, build_compressor, code_width,
                            sparsify, desparsify, topk_count, qsgd_quantize, qsgd_dequantize)
from ewdml.compress.rng import philox4x32, uniform_for_indices, uniform_like


def restated_qsgd(t, s, u):
    """The arithmetic of src/Compresssor/qsgd.py:17-28 with the uniform tensor injected."""
    norm = t.norm().flatten()
    lf = s / norm * t.abs()
    prev = lf.floor()
    lvl = prev + (u < (lf - prev))
    return lvl * t.sign(), norm


def test_qsgd_matches_restated_reference_math():
    torch.manual_seed(0)
    t = torch.randn(7, 33)
    u = torch.rand_like(t)
    c = QSGDCompressor(128, pack=False)
    q, norm = c.compress(t, uniform=u)
    q_ref, norm_ref = restated_qsgd(t, 128, u)
    assert torch.equal(q, q_ref)
    assert torch.equal(norm, norm_ref)
    # both decompress conventions
    d1 = c.decompress((q, norm))
    d2 = c.decompress(q, norm)
    assert torch.equal(d1, d2)
    assert torch.allclose(d1, norm_ref / 128 * q_ref)


@pytest.mark.parametrize("s", [4, 16, 64, 128, 256])
def test_qsgd_levels_in_range_and_error_bound(s):
    torch.manual_seed(1)
    t = torch.randn(4096) * 3
    c = QSGDCompressor(s, seed=3, pack=False)
    q, norm = c.compress(t)
    assert q.abs().max() <= s and q.abs().min() >= 0
    assert torch.equal(q, q.round())
    d = c.decompress(q, norm)
    # each element is off by at most one quantisation step
    assert (d - t).abs().max() <= norm / s + 1e-6


def test_qsgd_unbiased():
    torch.manual_seed(2)
    t = torch.randn(256)
    c = QSGDCompressor(8, seed=5, pack=False)
    acc = torch.zeros_like(t)
    n = 400
    for step in range(n):
        c.step = step
        acc += c.decompress(*c.compress(t))
    err = (acc / n - t).abs().max().item()
    step_size = (t.norm() / 8).item()
    assert err < 4 * step_size / math.sqrt(n) * 1.5


def test_qsgd_zero_norm_and_packing_dtype():
    c = QSGDCompressor(128)
    q, norm = c.compress(torch.zeros(10))
    assert norm.item() == 0 and q.abs().sum() == 0 and not torch.isnan(c.decompress(q, norm)).any()
    assert QSGDCompressor(127).compress(torch.randn(5))[0].dtype == torch.int8
    assert QSGDCompressor(128).compress(torch.randn(5))[0].dtype == torch.int16
    assert code_width(128, sparse=True) == 8 and code_width(128, sparse=False) == 9
    assert code_width(4, sparse=True) == 3 and code_width(256, sparse=True) == 9


def test_qsgd_single_nonzero_hits_top_level():
    t = torch.zeros(8); t[3] = -2.5
    q, norm = QSGDCompressor(64, pack=False).compress(t)
    assert q[3] == -64 and norm.item() == pytest.approx(2.5)


def test_qsgd_deterministic_counter_rng():
    t = torch.randn(1000)
    a = QSGDCompressor(128, seed=11); b = QSGDCompressor(128, seed=11)
    assert torch.equal(a.compress(t)[0], b.compress(t)[0])
    b.step = 1
    assert not torch.equal(a.compress(t)[0], b.compress(t)[0])
    # offset = global element index: quantising a slice equals slicing the quantised whole
    whole = uniform_like(torch.empty(100), 11, 0)
    part = uniform_like(torch.empty(40), 11, 0, offset=60)
    assert torch.equal(whole[60:], part)


def test_philox_known_answer():
    # Random123 known-answer vectors for philox4x32-10
    r = philox4x32([0], [0], [0], [0], 0, 0)
    assert [int(x[0]) for x in r] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    r = philox4x32([0xffffffff], [0xffffffff], [0xffffffff], [0xffffffff], 0xffffffff, 0xffffffff)
    assert [int(x[0]) for x in r] == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    u = uniform_for_indices(torch.arange(10000), 1, 2)
    assert 0 <= u.min() and u.max() < 1 and abs(u.mean().item() - 0.5) < 0.02


@pytest.mark.parametrize("ratio,n", [(0.4, 12), (0.5, 1000), (0.01, 5000), (0.001, 100), (1.0, 17)])
def test_topk_exact_k_and_roundtrip(ratio, n):
    torch.manual_seed(3)
    t = torch.randn(n)
    c = TopKCompressor(ratio)
    (vals, idx), ctx = c.compress(t.view(-1, 1) if n % 2 else t)
    k = max(1, int(n * ratio))
    assert vals.numel() == k == topk_count(n, ratio) and idx.numel() == k
    # same *set* as torch.topk (reference TopK.py:8)
    ref_idx = torch.topk(t.abs(), k, sorted=False).indices
    assert set(idx.tolist()) == set(ref_idx.tolist())
    assert torch.equal(vals, t[idx.long()])
    d = c.decompress((vals, idx), ctx)
    assert d.shape == ctx[1]
    assert torch.equal(d.flatten()[idx.long()], vals) and int((d != 0).sum()) <= k


def test_topk_ties_lowest_index_first():
    t = torch.tensor([1.0, -2.0, 2.0, 2.0, 0.5, -2.0])
    vals, idx = sparsify(t, 0.5)          # k = 3 of four |2.0| entries
    assert idx.tolist() == [1, 2, 3]
    assert desparsify((vals, idx), 6).tolist() == [0, -2, 2, 2, 0, 0]


def test_topk_qsgd_order_and_norm():
    torch.manual_seed(4)
    t = torch.randn(50, 20)
    u = torch.rand(1000)
    c = TopKQSGDCompressor(0.1, 128)
    (q, idx), (norm, numel, shape) = c.compress(t, uniform=u)
    kept = t.flatten()[idx.long()]
    assert norm.item() == pytest.approx(kept.norm().item(), rel=1e-6)     # norm of the kept values
    sparse = torch.zeros(1000); sparse[idx.long()] = kept
    q_ref, _ = restated_qsgd(sparse, 128, u)                               # topk -> densify -> qsgd
    dense_q = torch.zeros(1000); dense_q[idx.long()] = q.float()
    assert torch.equal(dense_q, q_ref)
    d = c.decompress((q, idx), (norm, numel, shape))
    assert d.shape == t.shape and int((d != 0).sum()) <= 100


def test_horovod_flavour_and_namespace():
    t = torch.randn(300)
    q, ctx = HorovodQSGDCompressor.compress(t)
    d = HorovodQSGDCompressor.decompress(q, ctx)
    assert (d - t).abs().max() <= ctx / 128 + 1e-6
    assert Compression.qsgd is HorovodQSGDCompressor
    p, c = Compression.fp16().compress(t)
    assert p.dtype == torch.float16 and Compression.fp16().decompress(p, c).dtype == torch.float32
    assert torch.equal(NoneCompressor().roundtrip(t), t)


def test_payload_bytes_beat_reference_ratios():
    """method 4 = 4x, method 5 (k=0.4) = 5x vs fp32."""
    torch.manual_seed(5)
    t = torch.randn(400000) * 0.01
    dense = 4 * t.numel()
    c4 = build_compressor("qsgd", 128)
    assert dense / c4.payload_nbytes(*c4.compress(t)) >= 4.0
    c5 = build_compressor("topk-qsgd", 128, 0.4)
    assert dense / c5.payload_nbytes(*c5.compress(t)) >= 5.0
