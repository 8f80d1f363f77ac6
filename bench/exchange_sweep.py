#!/usr/bin/env python
"""Compressed-exchange sweep (BASELINE.json config 5): gradient sizes 1 MB - 1 GB, k = 0.1 - 10 %, QSGD s = 4 - 256,
at the launched world size, against the library path (torch.topk + QSGD torch ops + NCCL all_gather, and NCCL
all_reduce for the dense case).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/exchange_sweep.py [--quick]

Per point: device time (CUDA events, max over ranks, 3 warm-ups, median of `--iters`) of K1 send + K2 receive/average,
payload bytes per peer, bytes over NVLink per rank, effective exchange rate (dense gradient bytes / time) and the
roofline fraction: max(2 HBM reads of the gradient at the measured copy bandwidth, (W-1) * payload / 770 GB/s) / time.
One JSON line per point on rank 0.
"""
import argparse, json, os, sys, statistics
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ewdml.parallel.native_exchange import ExchangePlan, Mailbox, make_codec
from ewdml.parallel.peer import PeerGroup
from ewdml.runtime import lib as L

HBM_GBS, NVLINK_GBS = 6557.8, 770.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--iters", type=int, default=7)
    ap.add_argument("--bucket", type=int, default=1 << 22, help="elements per bucket (segment)")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    pg = PeerGroup.from_env(dev)
    peaks = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    hbm = json.load(open(peaks))["hbm_gbs"] if os.path.exists(peaks) else HBM_GBS
    sizes_mb = [1, 16, 256] if a.quick else [1, 16, 64, 256, 1024]
    combos = [("topk-qsgd", 0.01, 128)] if a.quick else \
        [("topk-qsgd", r, 128) for r in (0.001, 0.01, 0.1)] + [("topk-qsgd", 0.01, s) for s in (4, 16, 256)] + \
        [("topk", 0.01, 128), ("qsgd", 1.0, 128), ("none", 1.0, 128)]

    def timeit2(send, recv):
        """median (K1 ms, K2 ms, total ms), each the max over ranks: K1 = select + encode + push, K2 = wait + decode + average."""
        rows = []
        for i in range(3 + a.iters):
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            e[0].record(); send(); e[1].record(); recv(); e[2].record(); torch.cuda.synchronize()
            t = torch.tensor([e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2]), e[0].elapsed_time(e[2])], device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if i >= 3:
                rows.append(t.tolist())
        return [statistics.median(r[j] for r in rows) for j in range(3)]

    def timeit(fn):
        ts = []
        for i in range(3 + a.iters):
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if i >= 3:
                ts.append(float(t))
        return statistics.median(ts)

    for mb in sizes_mb:
        n = mb * (1 << 20) // 4
        offs, nums, o = [], [], 0
        while o < n:
            m = min(a.bucket, n - o); offs.append(o); nums.append(m); o += (m + 31) // 32 * 32
        g = torch.randn(o, device=dev) * 0.01
        out = torch.empty(o, device=dev)
        for comp, ratio, s in combos:
            if comp == "none" and mb * world * 2 > 6000:
                continue
            codec = make_codec(comp, s, ratio)
            plan = ExchangePlan(offs, nums, codec)
            mb_ = Mailbox(pg.alloc(f"sw{mb}{comp}{ratio}{s}", Mailbox.nbytes(world, plan.slot_bytes)), world, plan.slot_bytes)
            state = torch.zeros(4, dtype=torch.int32, device=dev)
            lib = L.load()
            import ctypes as C

            def k1():
                L.check(lib.ewdml_state_advance(C.c_void_p(state.data_ptr()), 1, ExchangePlan._stream()))
                plan.send_state(g, [(mb_.slot(r, 0, rank), mb_.flag(r, 0, rank)) for r in range(world)],
                                [(mb_.slot(r, 1, rank), mb_.flag(r, 1, rank)) for r in range(world)], state, 1, rank,
                                mb_.mc_slot(0, rank), mb_.mc_slot(1, rank))

            def k2():
                plan.recv_apply_state([(mb_.slot(rank, 0, r), mb_.flag(rank, 0, r)) for r in range(world)],
                                      [(mb_.slot(rank, 1, r), mb_.flag(rank, 1, r)) for r in range(world)], state,
                                      float(world), L.OUT_GRAD, out=out)
            t_k1, t_k2, t_ours = timeit2(k1, k2)
            st = plan.stats()
            payload = st["payload_bytes"]
            multicast = bool(mb_.mc_slot(0, rank))
            # library path
            k = max(1, int(n * ratio))
            if comp in ("topk-qsgd", "topk"):
                vals_all = torch.empty(world * k, device=dev); idx_all = torch.empty(world * k, dtype=torch.int64, device=dev)

                def lib_path():
                    _, idx = torch.topk(g[:n].abs(), k, sorted=False)
                    v = g[:n][idx]
                    if comp == "topk-qsgd":
                        norm = v.norm(); lf = s / norm * v.abs(); pv = lf.floor()
                        v = (pv + (torch.rand_like(v) < (lf - pv))) * v.sign() * (norm / s)
                    if world > 1:
                        dist.all_gather_into_tensor(vals_all, v); dist.all_gather_into_tensor(idx_all, idx)
                    else:
                        vals_all.copy_(v); idx_all.copy_(idx)
                    out[:n].zero_(); out[:n].index_add_(0, idx_all, vals_all); out[:n].div_(world)
            else:
                def lib_path():
                    out[:n].copy_(g[:n])
                    if world > 1:
                        dist.all_reduce(out[:n])
                    out[:n].div_(world)
            t_lib = timeit(lib_path) if mb <= 256 or comp == "none" else None
            # each kernel chain against ITS OWN bytes: K1 reads the gradient twice (band pass + encode pass; once for the dense and
            # QSGD-only codecs); on the wire an allgather delivers (W-1) payloads INTO every rank whatever the sender does --
            # multicast only cuts the sender's egress from (W-1) payloads to one -- so the link bound is the ingress; K2 reads
            # the W payloads and writes the dense average
            reads = 1 if comp in ("none", "qsgd") else 2
            wire = (world - 1) * payload
            roof_k1 = max(reads * n * 4 / (hbm * 1e9), wire / (NVLINK_GBS * 1e9)) * 1e3
            roof_k2 = (world * payload + n * 4) / (hbm * 1e9) * 1e3
            t_roof = roof_k1 + roof_k2
            if rank == 0:
                print(json.dumps({"gradient_mb": mb, "world": world, "compressor": comp, "ratio": ratio, "levels": s,
                                  "multicast": multicast, "k1_ms": round(t_k1, 4), "k2_ms": round(t_k2, 4),
                                  "k1_roofline_ms": round(roof_k1, 4), "k2_roofline_ms": round(roof_k2, 4),
                                  "k1_fraction": round(roof_k1 / t_k1, 3), "k2_fraction": round(roof_k2 / t_k2, 3),
                                  "ms": round(t_ours, 4), "library_ms": None if t_lib is None else round(t_lib, 4),
                                  "speedup_vs_library": None if t_lib is None else round(t_lib / t_ours, 2),
                                  "payload_bytes": int(payload), "nvlink_bytes_per_rank": int((world - 1) * payload),
                                  "effective_GBps": round(n * 4 / t_ours / 1e6, 1), "roofline_ms": round(t_roof, 4),
                                  "roofline_fraction": round(t_roof / t_ours, 3)}), flush=True)
            plan.close()
            del mb_
    if world > 1:
        dist.barrier()
    pg.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
