"""GPU numerics tests of the exchange kernels against the plain-PyTorch fp32 oracle.

Multi-rank behaviour is emulated on ONE GPU by :class:`LocalPeerGroup`: all senders run first, then
all receivers, on one stream (kernels that wait on each other are never co-scheduled on one GPU).
Real multi-GPU runs live in tests/test_multigpu.py.
"""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from ewdml.compress.topk import topk_count
from ewdml.optim import SGD
from ewdml.parallel.flat import Segment
from ewdml.parallel.host_codec import HostCodec


def _lib():
    from ewdml.runtime import lib
    return lib


def _segments(sizes):
    segs, off = [], 0
    for i, n in enumerate(sizes):
        segs.append(Segment(f"p{i}", off, n, torch.Size([n]), False))
        off = (off + n + 31) // 32 * 32
    return segs, max(off, 32)


class Emu:
    """W emulated ranks exchanging through local mailboxes."""

    def __init__(self, sizes, compressor, W=1, levels=128, ratio=0.01, seed=7):
        from ewdml.parallel.native_exchange import ExchangePlan, Mailbox, make_codec
        from ewdml.parallel.peer import LocalPeerGroup
        self.dev = torch.device("cuda", 0)
        self.W, self.seed, self.compressor, self.levels, self.ratio = W, seed, compressor, levels, ratio
        self.segs, self.total = _segments(sizes)
        self.codec = make_codec(compressor, levels, ratio)
        offs, nums = [s.offset for s in self.segs], [s.numel for s in self.segs]
        self.plans = [ExchangePlan(offs, nums, self.codec) for _ in range(W)]
        pg = LocalPeerGroup(W, self.dev)
        syms = pg.alloc("mb", Mailbox.nbytes(W, self.plans[0].slot_bytes))
        self.mbs = [Mailbox(s, W, self.plans[0].slot_bytes) for s in syms]
        self.q = 0

    def grads(self, gen_seed=0, scale=1.0, kind="randn"):
        g = torch.Generator(device="cpu").manual_seed(gen_seed)
        out = []
        for r in range(self.W):
            t = torch.zeros(self.total)
            for s in self.segs:
                if kind == "randn":
                    v = torch.randn(s.numel, generator=g) * scale
                elif kind == "heavy":
                    v = torch.randn(s.numel, generator=g) ** 3 * scale
                t[s.offset:s.offset + s.numel] = v
            out.append(t.to(self.dev))
        return out

    def exchange(self, grads, step=0, out_mode=None, **kw):
        L = _lib()
        self.q += 1
        q, par = self.q, self.q & 1
        for r in range(self.W):
            self.plans[r].send(grads[r], [(self.mbs[r].slot(d, par, r), self.mbs[r].flag(d, par, r)) for d in range(self.W)],
                               q, self.seed, step, r)
        outs = []
        for r in range(self.W):
            srcs = [(self.mbs[r].slot(r, par, s), self.mbs[r].flag(r, par, s)) for s in range(self.W)]
            if out_mode is None:
                out = torch.full((self.total,), float("nan"), device=self.dev)
                self.plans[r].recv_apply(srcs, q, float(self.W), L.OUT_GRAD, out=out)
                outs.append(out)
            else:
                self.plans[r].recv_apply(srcs, q, float(self.W), out_mode, **kw["kw"][r])
        torch.cuda.synchronize()
        return outs

    def oracle(self, grads, step=0):
        acc = torch.zeros(self.total)
        for r in range(self.W):
            codec = HostCodec(self.segs, self.compressor, self.levels, self.ratio, seed=self.seed, rank=r,
                              precise_norm=True)
            codec.decode(codec.encode(grads[r].cpu(), step), out=acc, accumulate=True, total=self.total)
        return acc / self.W

    def valid_mask(self):
        m = torch.zeros(self.total, dtype=torch.bool)
        for s in self.segs:
            m[s.offset:s.offset + s.numel] = True
        return m


SIZES = [10, 500, 8192, 8193, 70001, 1200000]


def _assert_close_quantised(out, ref, grads, emu, max_bad_frac=2e-4):
    """QSGD results may differ in a handful of elements where u is within 1 ulp of the rounding
    fraction (the kernel's norm is summed in a different order than torch.norm)."""
    m = emu.valid_mask()
    o, r = out.cpu()[m], ref[m]
    diff = (o - r).abs()
    tol = 1e-5 * r.abs().max().clamp_min(1e-30)
    bad = diff > tol
    assert bad.float().mean().item() <= max_bad_frac, f"{int(bad.sum())} of {bad.numel()} elements differ"
    # the few that differ are off by at most one quantisation step of the largest-norm segment
    step = max(float(g.abs().max()) for g in grads) * 4 / emu.levels * 64
    assert diff.max().item() <= step


@pytest.mark.parametrize("W", [1, 3])
def test_dense_mode_is_exact_average(W):
    emu = Emu(SIZES, "none", W)
    g = emu.grads(1)
    outs = emu.exchange(g)
    ref = torch.stack([x.cpu() for x in g]).sum(0)
    # rank-order accumulation then division, like the oracle
    acc = torch.zeros(emu.total)
    for x in g:
        acc += x.cpu()
    ref = acc / W
    m = emu.valid_mask()
    for o in outs:
        assert torch.equal(o.cpu()[m], ref[m])


@pytest.mark.parametrize("ratio,W", [(0.01, 1), (0.01, 3), (0.4, 2), (0.001, 1), (1.0, 1)])
def test_topk_exact_selection_matches_oracle(ratio, W):
    emu = Emu(SIZES, "topk", W, ratio=ratio)
    g = emu.grads(2)
    outs = emu.exchange(g)
    ref = emu.oracle(g)
    m = emu.valid_mask()
    for o in outs:
        assert torch.equal(o.cpu()[m], ref[m])
    # exact k per segment: #(key > thr) + ties taken == k
    info = emu.plans[0].seginfo()
    for s, si in zip(emu.segs, info):
        k = topk_count(s.numel, ratio)
        keys = g[0][s.offset:s.offset + s.numel].abs().cpu().view(torch.int32).long()
        thr = si["thr_key"] - (1 << 32) if si["thr_key"] >= (1 << 31) else si["thr_key"]     # signed: -1 = select all
        n_gt = int((keys > thr).sum())
        n_eq = int((keys == thr).sum())
        assert n_gt + min(si["need_eq"], n_eq) == k
    st = emu.plans[0].stats()
    assert st["entries"] == sum(topk_count(s.numel, ratio) for s in emu.segs) and st["error"] == 0


@pytest.mark.parametrize("levels", [4, 128, 256])
@pytest.mark.parametrize("W", [1, 2])
def test_qsgd_dense_matches_oracle(levels, W):
    emu = Emu(SIZES, "qsgd", W, levels=levels)
    g = emu.grads(3, scale=0.02)
    outs = emu.exchange(g, step=5)
    ref = emu.oracle(g, step=5)
    for o in outs:
        _assert_close_quantised(o, ref, g, emu)
    # norms agree with torch to fp32 rounding
    info = emu.plans[0].seginfo()
    for s, si in zip(emu.segs, info):
        n_ref = g[0][s.offset:s.offset + s.numel].double().norm().item()
        assert si["norm"] == pytest.approx(n_ref, rel=2e-6)


@pytest.mark.parametrize("ratio,levels,W,kind", [(0.01, 128, 1, "randn"), (0.01, 128, 3, "heavy"), (0.4, 128, 2, "randn"),
                                                 (0.1, 16, 2, "randn"), (0.05, 256, 1, "heavy")])
def test_topk_qsgd_matches_oracle(ratio, levels, W, kind):
    emu = Emu(SIZES, "topk-qsgd", W, levels=levels, ratio=ratio)
    g = emu.grads(4, scale=0.1, kind=kind)
    for step in (0, 1, 2):                       # three exchanges: parity flip + counter reset
        outs = emu.exchange(g, step=step)
        ref = emu.oracle(g, step=step)
        for o in outs:
            _assert_close_quantised(o, ref, g, emu)
    assert all(si["fallback"] == 0 for si in emu.plans[0].seginfo())


def test_degenerate_segments_zero_constant_ties():
    sizes = [4096, 50000, 300, 33]
    emu = Emu(sizes, "topk-qsgd", 1, ratio=0.1)
    g = emu.grads(5)
    s0, s1, s2, s3 = emu.segs
    g[0][s0.offset:s0.offset + s0.numel] = 0.0                     # all zeros: norm 0 -> nothing sent
    g[0][s1.offset:s1.offset + s1.numel] = 0.5                     # constant: every element ties
    g[0][s2.offset:s2.offset + 150] = 0.0                          # half zeros
    out = emu.exchange(g)[0].cpu()
    assert not torch.isnan(out[emu.valid_mask()]).any()
    assert out[s0.offset:s0.offset + s0.numel].abs().sum() == 0
    c = out[s1.offset:s1.offset + s1.numel]
    k1 = topk_count(s1.numel, 0.1)
    assert int((c != 0).sum()) <= k1                                # exactly k ties taken, minus level-0 drops
    info = emu.plans[0].seginfo()
    assert info[1]["need_eq"] == k1 and info[1]["norm"] == pytest.approx(0.5 * math.sqrt(k1), rel=1e-5)   # genuine tie
    # topk (raw values) on the constant segment keeps exactly k elements
    emu2 = Emu(sizes, "topk", 1, ratio=0.1)
    out2 = emu2.exchange(g)[0].cpu()
    assert int((out2[s1.offset:s1.offset + s1.numel] != 0).sum()) == k1


def test_payload_bytes_and_compression_ratio():
    n = 2_000_000
    emu = Emu([n], "topk-qsgd", 1, ratio=0.01)
    g = emu.grads(6)
    emu.exchange(g)
    st = emu.plans[0].stats()
    k = topk_count(n, 0.01)
    assert st["entries"] <= k and st["entries"] > 0.5 * k
    # 2 B index + 1 B code per entry + small per-tile overhead
    assert st["data_bytes"] <= 3 * k + 48 * emu.plans[0].ntiles
    assert 4 * n / st["payload_bytes"] > 100        # > 100x smaller than the dense fp32 gradient
    emu4 = Emu([n], "qsgd", 1)
    g4 = emu4.grads(6)
    emu4.exchange(g4)
    assert 4 * n / emu4.plans[0].stats()["payload_bytes"] > 8     # reference accounting claims 4x


@pytest.mark.parametrize("kw", [dict(momentum=0.9), dict(momentum=0.9, nesterov=True, weight_decay=1e-3), dict(momentum=0.0)])
def test_recv_apply_sgd_matches_reference_optimizer(kw):
    L = _lib()
    from ewdml.config import TrainConfig
    from ewdml.parallel.native_exchange import make_sgd
    emu = Emu([5000, 40000], "none", 2)
    cfg = TrainConfig(lr=0.05, **kw)
    dev = emu.dev
    torch.manual_seed(0)
    p0 = torch.randn(emu.total)
    params = [p0.clone().to(dev) for _ in range(2)]
    moms = [torch.zeros(emu.total, device=dev) for _ in range(2)]
    shadows = [torch.zeros(emu.total, dtype=torch.bfloat16, device=dev) for _ in range(2)]
    ref_p = p0.clone().requires_grad_()
    opt = SGD([ref_p], lr=cfg.lr, momentum=cfg.momentum, weight_decay=cfg.weight_decay, nesterov=cfg.nesterov)
    for step in range(3):
        g = emu.grads(10 + step)
        kws = [dict(sgd=make_sgd(cfg, step == 0), param=params[r], momentum=moms[r], param_bf16=shadows[r]) for r in range(2)]
        emu.exchange(g, out_mode=L.OUT_SGD, kw=kws)
        avg = (g[0].cpu() + g[1].cpu()) / 2
        opt.step(grads=[avg])
    m = emu.valid_mask()
    for r in range(2):
        assert torch.allclose(params[r].cpu()[m], ref_p.detach()[m], rtol=1e-5, atol=1e-6)
        assert torch.equal(shadows[r].cpu()[m], params[r].cpu()[m].to(torch.bfloat16))


def test_local_sgd_apply_kernel():
    import ctypes as C
    L = _lib()
    lib = L.load()
    from ewdml.config import TrainConfig
    from ewdml.parallel.native_exchange import make_sgd
    dev = torch.device("cuda", 0)
    n = 1 << 20
    torch.manual_seed(1)
    p, g = torch.randn(n), torch.randn(n)
    dp, dg, dm = p.to(dev), g.to(dev), torch.zeros(n, device=dev)
    cfg = TrainConfig(lr=0.1, momentum=0.9)
    ref = p.clone().requires_grad_()
    opt = SGD([ref], lr=0.1, momentum=0.9)
    for step in range(2):
        sgd = make_sgd(cfg, step == 0)
        L.check(lib.ewdml_sgd_apply(C.c_void_p(dg.data_ptr()), C.c_void_p(dp.data_ptr()), C.c_void_p(dm.data_ptr()), None,
                                    n, C.byref(sgd), 1.0, C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        opt.step(grads=[g])
    torch.cuda.synchronize()
    assert torch.allclose(dp.cpu(), ref.detach(), rtol=1e-5, atol=1e-6)


def test_flag_timeout_sets_error_word():
    import ctypes as C
    L = _lib()
    lib = L.load()
    emu = Emu([1000], "none", 1)
    lib.ewdml_set_spin_limit(2000)
    try:
        out = torch.zeros(emu.total, device=emu.dev)
        # nobody ever sends exchange number 5
        emu.plans[0].recv_apply([(emu.mbs[0].slot(0, 1, 0), emu.mbs[0].flag(0, 1, 0))], 5, 1.0, L.OUT_GRAD, out=out)
        err = C.c_uint32(0)
        L.check(lib.ewdml_get_device_error(C.byref(err), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        assert err.value == (L_ERR_TIMEOUT & 0xFFFFFFFF)
    finally:
        lib.ewdml_set_spin_limit(0)


L_ERR_TIMEOUT = -3


def test_single_rank_allreduce_and_counters():
    import ctypes as C
    L = _lib()
    lib = L.load()
    from ewdml.parallel.peer import LocalPeerGroup
    dev = torch.device("cuda", 0)
    pg = LocalPeerGroup(1, dev)
    n = 1 << 16
    buf = pg.alloc("g", 2 * n * 4)[0]
    flags = pg.alloc("f", 4096)[0]
    x = buf.view(torch.float32, 0, n)
    x.copy_(torch.arange(n, dtype=torch.float32))
    comm = L.Comm()
    comm.rank, comm.world = 0, 1
    comm.bufs[0] = buf.ptrs[0]
    comm.peer_flags[0] = flags.ptrs[0]
    comm.flags = flags.ptrs[0]
    before = lib.ewdml_counter_get(L.CTR_LAUNCHES)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for q, algo in ((1, L.AR_TWOSHOT), (2, L.AR_ONESHOT)):
        L.check(lib.ewdml_allreduce(C.byref(comm), n, 0.5, algo, q, C.c_void_p(flags.ptrs[0] + 2048), st))
    L.check(lib.ewdml_barrier(C.byref(comm), 3, st))
    torch.cuda.synchronize()
    assert torch.equal(x.cpu(), torch.arange(n, dtype=torch.float32) * 0.25)
    assert lib.ewdml_counter_get(L.CTR_LAUNCHES) - before == 4
