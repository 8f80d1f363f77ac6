"""Collective engine (N3) on one GPU.

``world`` emulated ranks share the device; every operation is driven as PUT on all ranks, then WAIT on all
ranks, so no kernel ever waits for a later launch.  Results are compared with plain torch on the same
inputs.  Several rounds per operation exercise both slot parities and the acknowledgement flags.  The
single-kernel (PHASE_BOTH) form, message pipelining in slot-sized pieces and the two-shot allreduce are
covered with world == 1 here and with real peers in tests/test_multigpu.py.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch.device("cuda", 0)


def _engines(dev, world, slot=1 << 16):
    from ewdml.parallel.collectives import CollectiveEngine
    from ewdml.parallel.peer import LocalPeerGroup
    return CollectiveEngine.create_local(LocalPeerGroup(world, dev), slot_payload=slot)


def _device_error():
    from ewdml.runtime import lib as L
    err = C.c_uint32(0)
    L.check(L.load().ewdml_get_device_error(C.byref(err), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return err.value


def _both(engs, fn):
    from ewdml.runtime import lib as L
    for e in engs:
        fn(e, L.PHASE_PUT)
    for e in engs:
        fn(e, L.PHASE_WAIT)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_gather_type_collectives(dev, world):
    engs = _engines(dev, world)
    n = 1000 * 4            # floats per piece, 16000 B
    g = torch.Generator(device="cpu").manual_seed(world)
    for rnd in range(4):
        x = [torch.randn(n, generator=g).to(dev) for _ in range(world)]
        # allgather
        out = [torch.zeros(world * n, device=dev) for _ in range(world)]
        _both(engs, lambda e, ph: e.all_gather(x[e.rank], out[e.rank], phases=ph))
        ref = torch.cat(x)
        assert all(torch.equal(o, ref) for o in out)
        # broadcast from a moving root
        root = rnd % world
        b = [x[r].clone() for r in range(world)]
        _both(engs, lambda e, ph: e.broadcast(b[e.rank], root, phases=ph))
        assert all(torch.equal(t, x[root]) for t in b)
        # gather to root
        gd = torch.zeros(world * n, device=dev)
        _both(engs, lambda e, ph: e.gather(x[e.rank], gd if e.rank == root else None, root, phases=ph))
        assert torch.equal(gd, ref)
        # scatter from root
        sd = [torch.zeros(n, device=dev) for _ in range(world)]
        _both(engs, lambda e, ph: e.scatter(ref if e.rank == root else None, sd[e.rank], root, phases=ph))
        assert all(torch.equal(sd[r], x[r]) for r in range(world))
        # alltoall
        big = [torch.randn(world * n, generator=g).to(dev) for _ in range(world)]
        a2a = [torch.zeros(world * n, device=dev) for _ in range(world)]
        _both(engs, lambda e, ph: e.all_to_all(big[e.rank], a2a[e.rank], phases=ph))
        for r in range(world):
            exp = torch.cat([big[s][r * n:(r + 1) * n] for s in range(world)])
            assert torch.equal(a2a[r], exp)
        _both(engs, lambda e, ph: e.barrier(phases=ph))
    torch.cuda.synchronize()
    assert _device_error() == 0


@pytest.mark.parametrize("world", [2, 8])
def test_reduce_type_collectives(dev, world):
    engs = _engines(dev, world)
    n = 2048
    g = torch.Generator(device="cpu").manual_seed(10 + world)
    for rnd in range(3):
        x = [torch.randn(n, generator=g).to(dev) for _ in range(world)]
        ordered = x[0].clone()
        for r in range(1, world):
            ordered += x[r]                          # the kernel sums in rank order
        # allreduce (one-shot), average
        ar = [t.clone() for t in x]
        _both(engs, lambda e, ph: e.all_reduce(ar[e.rank], scale=1.0 / world, phases=ph))
        for t in ar:
            assert torch.allclose(t, ordered / world, rtol=1e-6, atol=1e-7)
        assert all(torch.equal(ar[0], t) for t in ar)          # bit-identical on every rank
        # reduce to root
        root = (rnd + 1) % world
        rd = torch.zeros(n, device=dev)
        _both(engs, lambda e, ph: e.reduce(x[e.rank], rd if e.rank == root else None, root, phases=ph))
        assert torch.allclose(rd, ordered, rtol=1e-6, atol=1e-6)
        # reduce_scatter
        big = [torch.randn(world * n, generator=g).to(dev) for _ in range(world)]
        rs = [torch.zeros(n, device=dev) for _ in range(world)]
        _both(engs, lambda e, ph: e.reduce_scatter(big[e.rank], rs[e.rank], scale=0.5, phases=ph))
        tot = big[0].clone()
        for r in range(1, world):
            tot += big[r]
        for r in range(world):
            assert torch.allclose(rs[r], 0.5 * tot[r * n:(r + 1) * n], rtol=1e-6, atol=1e-6)
    # odd sizes go through the padded helper
    odd = [torch.full((37,), float(r + 1), device=dev) for r in range(world)]
    _both(engs, lambda e, ph: e.all_reduce(odd[e.rank], phases=ph))
    assert all(torch.equal(t, torch.full((37,), world * (world + 1) / 2.0, device=dev)) for t in odd)
    torch.cuda.synchronize()
    assert _device_error() == 0


def test_send_recv_in_order_with_tags(dev):
    engs = _engines(dev, 3)
    msgs = [torch.arange(256, dtype=torch.float32, device=dev) + 1000 * i for i in range(5)]
    got = [torch.zeros(256, device=dev) for _ in range(5)]
    # rank 0 -> rank 2, two messages in flight (both parities), then the receiver drains them; repeated
    for i in (0, 1):
        engs[0].send(msgs[i], peer=2, tag=7 + i)
    for i in (0, 1):
        engs[2].recv(got[i], peer=0, tag=7 + i)
    for i in (2, 3):
        engs[0].send(msgs[i], peer=2, tag=7 + i)
    engs[1].send(msgs[4], peer=2, tag=99)
    for i in (2, 3):
        engs[2].recv(got[i], peer=0, tag=7 + i)
    engs[2].recv(got[4], peer=1, tag=99)
    torch.cuda.synchronize()
    assert all(torch.equal(a, b) for a, b in zip(msgs, got))
    assert _device_error() == 0
    assert engs[0].calls("send") == 4 and engs[2].calls("recv") == 5


def test_single_kernel_form_pipelining_and_twoshot_world1(dev):
    from ewdml.parallel.collectives import CollectiveEngine
    from ewdml.parallel.peer import PeerGroup
    pg = PeerGroup(0, 1, dev)
    eng = CollectiveEngine.create(pg, slot_payload=4096)       # 1024 floats per slot -> many pieces
    x = torch.randn(10000, device=dev)
    y = torch.zeros(10000, device=dev)
    eng.all_gather(x, y)
    assert torch.equal(x, y)
    z = x.clone()
    eng.all_reduce(z, scale=2.0)
    assert torch.equal(z, 2.0 * x)
    assert eng.select_algo("allreduce", 1 << 30) == "oneshot"   # world <= 2
    eng.set_algo("allreduce", "twoshot")
    assert eng.select_algo("allreduce", 64) == "twoshot"
    w = x.clone()
    eng.all_reduce(w, scale=0.5)                                # reduce_scatter + allgather pieces, in place
    assert torch.equal(w, 0.5 * x)
    req = eng.iall_reduce(w, scale=2.0)                         # non-blocking form on the side stream
    req.wait()
    torch.cuda.synchronize()
    assert req.test() and torch.equal(w, x)
    assert abs(eng.metric_average(3.5) - 3.5) < 1e-6
    eng.barrier()
    torch.cuda.synchronize()
    assert _device_error() == 0


def test_argument_errors_are_reported(dev):
    from ewdml.runtime import lib as L
    engs = _engines(dev, 2, slot=4096)
    big = torch.zeros(4096, device=dev)                          # 16 KiB > slot payload
    with pytest.raises(L.NativeError):
        engs[0].all_gather(big, torch.zeros(8192, device=dev), phases=L.PHASE_PUT)   # split phases need one piece
    with pytest.raises(TypeError):
        engs[0].all_reduce(torch.zeros(16, dtype=torch.float16, device=dev))
    with pytest.raises(ValueError):
        engs[0].all_gather(torch.zeros(3, device=dev), torch.zeros(6, device=dev))


def test_engine_on_symmetric_heap(dev):
    """Window and flags carved out of one pre-mapped region by the deterministic sub-allocator."""
    from ewdml.parallel.collectives import CollectiveEngine
    from ewdml.parallel.peer import PeerGroup, SymmetricHeap
    pg = PeerGroup(0, 1, dev)
    heap = SymmetricHeap(pg, 8 << 20)
    scratch = heap.alloc("scratch", 1 << 20)
    eng = CollectiveEngine.create(pg, slot_payload=1 << 16, heap=heap)
    used = heap.used
    heap.free(scratch)
    assert heap.used == used - (1 << 20)
    x = torch.randn(40000, device=dev)
    y = x.clone()
    eng.all_reduce(y, scale=3.0)
    z = torch.zeros_like(x)
    eng.all_gather(x, z)
    torch.cuda.synchronize()
    assert torch.equal(y, 3.0 * x) and torch.equal(z, x)
    again = heap.alloc("again", 1 << 19)
    assert again.heap_offset == 0 and again.local.numel() == 1 << 19      # first fit: the freed block is reused
    assert _device_error() == 0
