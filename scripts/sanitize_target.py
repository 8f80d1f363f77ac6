"""Small, fast exercise of the communication kernels for compute-sanitizer:

    python scripts/sanitize_target.py                                     # must exit 0 on its own first
    compute-sanitizer --tool memcheck  python scripts/sanitize_target.py
    compute-sanitizer --tool racecheck python scripts/sanitize_target.py  # shared-memory hazards
    compute-sanitizer --tool synccheck python scripts/sanitize_target.py

Covers K1/K2 (top-k + QSGD, bitmap and index wire modes, dense), K4 (one rank), the collective engine (put + wait phases
of every operation, send/recv) and the fused SGD kernel, on emulated ranks of one GPU, with small tensors so the
instrumented run stays short.  Prints one JSON line.
"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from ewdml.parallel.collectives import CollectiveEngine
from ewdml.parallel.native_exchange import ExchangePlan, Mailbox, make_codec, make_sgd
from ewdml.parallel.peer import LocalPeerGroup
from ewdml.runtime import lib as L
from ewdml.config import TrainConfig


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    lib = L.load()
    W = 2
    sizes, offs, off = [1000, 70001, 9000], [], 0
    for n in sizes:
        offs.append(off)
        off = (off + n + 31) // 32 * 32
    total = off
    checks = {}
    cfg = TrainConfig(lr=0.1, momentum=0.9)
    for comp, ratio in (("topk-qsgd", 0.01), ("topk-qsgd", 0.3), ("qsgd", 1.0), ("topk", 0.05), ("none", 1.0)):
        codec = make_codec(comp, 128, ratio)
        plans = [ExchangePlan(offs, sizes, codec) for _ in range(W)]
        syms = LocalPeerGroup(W, dev).alloc("mb", Mailbox.nbytes(W, plans[0].slot_bytes))
        mbs = [Mailbox(s, W, plans[0].slot_bytes) for s in syms]
        grads = [torch.randn(total, device=dev) * 0.1 for _ in range(W)]
        param = [torch.zeros(total, device=dev) for _ in range(W)]
        mom = [torch.zeros(total, device=dev) for _ in range(W)]
        for q in (1, 2, 3):
            par = q & 1
            for r in range(W):
                plans[r].send(grads[r], [(mbs[d].slot(d, par, r), mbs[d].flag(d, par, r)) for d in range(W)], q, 7, q, r)
            for r in range(W):
                plans[r].recv_apply([(mbs[r].slot(r, par, s), mbs[r].flag(r, par, s)) for s in range(W)], q, float(W), L.OUT_SGD,
                                    make_sgd(cfg, q == 1), param[r], mom[r])
        torch.cuda.synchronize()
        checks[f"{comp}@{ratio}"] = bool(torch.equal(param[0], param[1]) and torch.isfinite(param[0]).all() and param[0].abs().sum() > 0)
    engs = CollectiveEngine.create_local(LocalPeerGroup(4, dev), slot_payload=1 << 14)
    n = 1024
    x = [torch.full((n,), float(r + 1), device=dev) for r in range(4)]
    ag = [torch.zeros(4 * n, device=dev) for _ in range(4)]
    a2a = [torch.zeros(4 * n, device=dev) for _ in range(4)]
    big = [torch.arange(4 * n, dtype=torch.float32, device=dev) + 100 * r for r in range(4)]
    rs = [torch.zeros(n, device=dev) for _ in range(4)]
    for rnd in range(3):
        for fn in (lambda e, ph: e.all_gather(x[e.rank], ag[e.rank], phases=ph),
                   lambda e, ph: e.all_reduce(x[e.rank].clone(), out=rs[e.rank], phases=ph),
                   lambda e, ph: e.all_to_all(big[e.rank], a2a[e.rank], phases=ph),
                   lambda e, ph: e.reduce_scatter(big[e.rank], rs[e.rank], phases=ph),
                   lambda e, ph: e.broadcast(ag[e.rank], rnd, phases=ph),
                   lambda e, ph: e.barrier(phases=ph)):
            for e in engs:
                fn(e, L.PHASE_PUT)
            for e in engs:
                fn(e, L.PHASE_WAIT)
    got = torch.zeros(n, device=dev)
    engs[0].send(x[0], 3, tag=1)
    engs[3].recv(got, 0, tag=1)
    torch.cuda.synchronize()
    checks["engine"] = bool(torch.equal(got, x[0]) and torch.equal(ag[1], ag[2]))
    g = torch.randn(4096, device=dev); p = torch.zeros(4096, device=dev); m = torch.zeros(4096, device=dev)
    sgd = make_sgd(cfg, True)
    L.check(lib.ewdml_sgd_apply(C.c_void_p(g.data_ptr()), C.c_void_p(p.data_ptr()), C.c_void_p(m.data_ptr()), None, 4096, C.byref(sgd),
                                1.0, C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    checks["sgd"] = bool(torch.allclose(p, -0.1 * g))
    err = C.c_uint32(0)
    L.check(lib.ewdml_get_device_error(C.byref(err), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    ok = all(checks.values()) and err.value == 0
    print(json.dumps({"ok": ok, "device_error": err.value, "checks": checks, "launches": L.launches()}))
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
