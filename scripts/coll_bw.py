"""Device-timed throughput of the collective engine's kernel on one GPU (world 1: put into the own window + consume)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ewdml.parallel.collectives import CollectiveEngine
from ewdml.parallel.peer import PeerGroup
dev = torch.device("cuda", 0)
pg = PeerGroup(0, 1, dev)
eng = CollectiveEngine.create(pg, slot_payload=64 << 20)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for mb in (1, 16, 64):
    t = torch.randn(mb << 18, device=dev)
    o = torch.empty_like(t)
    for op, fn in (("allreduce", lambda: eng.all_reduce(t, out=o, scale=0.5)), ("allgather", lambda: eng.all_gather(t, o))):
        ms = []
        for it in range(8):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            if it >= 3:
                ms.append(e0.elapsed_time(e1))
        m = sorted(ms)[len(ms) // 2]
        print(json.dumps({"op": op, "world": 1, "mbytes": mb, "ms": m, "GBps_moved": 4 * mb * 2 ** 20 / m / 1e6,
                          "note": "bytes moved = read src + write window + read window + write dst"}))
