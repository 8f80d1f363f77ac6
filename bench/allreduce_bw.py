#!/usr/bin/env python
"""K4 dense fused-scale allreduce: two-shot (peer loads + unicast stores) vs NVLS (multimem.ld_reduce + multimem.st) vs NCCL.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/allreduce_bw.py [--sizes-mb 1 16 64 256 1024]

Per size and algorithm: device time (CUDA events around the one kernel, barrier before, max over ranks, median of --iters
after 3 warm-ups), algorithm bandwidth (bytes / time), bus bandwidth (2 (W-1)/W bytes / time, NCCL's convention) and the
fraction of the link bound: bytes that must cross one GPU's NVLink in one direction / 770 GB/s measured peer-copy rate --
2 (W-1)/W * B for two-shot (W-1 peer shards in, W-1 result shards out share each direction with the peers' traffic),
B for NVLS (the switch reduces and replicates).  One JSON line per point on rank 0.
"""
import argparse, ctypes as C, json, os, statistics, sys
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ewdml.parallel.peer import PeerGroup
from ewdml.runtime import lib as L

NVLINK_GBS = 770.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes-mb", type=int, nargs="+", default=[1, 16, 64, 256, 1024])
    ap.add_argument("--iters", type=int, default=9)
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    pg = PeerGroup.from_env(dev)
    lib = L.load()
    flags = pg.alloc("arbw_flags", 4096)
    seq = [0]

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

    for mb in a.sizes_mb:
        n = mb * (1 << 20) // 4
        buf = pg.alloc(f"arbw_{mb}", n * 4)
        x = buf.view(torch.float32, 0, n)
        x.fill_(1.0)
        comm = L.Comm()
        comm.rank, comm.world = rank, world
        for r in range(world):
            comm.bufs[r] = buf.ptrs[r]
            comm.peer_flags[r] = flags.ptrs[r]
        comm.flags = flags.ptrs[rank]
        comm.mc_buf = buf.mc_ptr or None
        st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
        algos = [("twoshot", L.AR_TWOSHOT)] + ([("nvls", L.AR_NVLS)] if buf.mc_ptr else [])
        y = torch.ones(n, device=dev)
        for name, algo in algos + [("nccl", None)]:
            if algo is None:
                if world == 1:
                    continue
                fn = lambda: dist.all_reduce(y)
            else:
                def fn(algo=algo):
                    seq[0] += 1
                    L.check(lib.ewdml_allreduce(C.byref(comm), n, 1.0 / world, algo, seq[0], C.c_void_p(flags.ptrs[rank] + 2048), st()))
            ms = timeit(fn)
            B = n * 4
            cross = B if name == "nvls" else 2 * (world - 1) / world * B
            if rank == 0:
                print(json.dumps({"op": "allreduce", "algo": name, "world": world, "mb": mb, "ms": round(ms, 4),
                                  "algbw_GBps": round(B / ms / 1e6, 1), "busbw_GBps": round(2 * (world - 1) / world * B / ms / 1e6, 1),
                                  "link_bound_ms": round(cross / (NVLINK_GBS * 1e9) * 1e3, 4),
                                  "fraction_of_link_bound": round(cross / (NVLINK_GBS * 1e9) * 1e3 / ms, 3) if world > 1 else None}), flush=True)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        del x, buf
    err = C.c_uint32(0)
    L.check(lib.ewdml_get_device_error(C.byref(err), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    if rank == 0:
        print(json.dumps({"device_error": err.value, "heap": pg.mode, "multicast": bool(pg.multicast)}), flush=True)
    if world > 1:
        dist.barrier()
    pg.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
