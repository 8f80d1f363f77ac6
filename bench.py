#!/usr/bin/env python
"""Headline benchmark: images/s of one compressed-gradient data-parallel training step.

    python bench.py --gpus N --steps K --warmup W            # this framework
    python bench.py --impl reference --gpus N ...             # the unmodified reference (baseline/_ref)

Config (BASELINE.json config 2): VGG-11 (BatchNorm variant, the reference's "VGG11") on synthetic
CIFAR-10-shaped data, random-init weights, per-GPU batch 128 (the reference entry point's default
``--batch-size``), top-k 1 % + QSGD (s = 128) sparse allgather, bf16 compute over fp32 master weights,
weak scaling.  One process per GPU; for N > 1 the driver launches this file under
``torch.distributed.run``.

Two timed regions, both W warm-up steps then exactly K steps between barrier + synchronize:
``value`` uses a device-resident batch (CUDA events, max over ranks); ``e2e`` goes through the
public API with, every step, the pinned-host -> device copy of that step's batch and a device ->
host read of the loss.  A 256 MiB buffer is overwritten every step inside the timed region to
flush L2.  ``gpu_launches`` counts the launches of this repo's own kernels (libewdml.so) inside
the first timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)



def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=200)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", type=str, default="ewdml", choices=["ewdml", "reference", "nccl"])
    p.add_argument("--model", type=str, default="VGG11")
    p.add_argument("--batch-size", type=int, default=128, help="per-GPU batch")
    p.add_argument("--compressor", type=str, default="topk-qsgd")
    p.add_argument("--topk-ratio", type=float, default=0.01)
    p.add_argument("--qsgd-levels", type=int, default=128)
    p.add_argument("--exchange", type=str, default="allgather", choices=["allgather", "allreduce", "ps"])
    p.add_argument("--compute", type=str, default=os.environ.get("EWDML_BENCH_COMPUTE", "native"),
                   choices=["torch", "native"])
    p.add_argument("--dtype", type=str, default="bf16")
    p.add_argument("--bucket-size", type=int, default=0, help="bytes per exchange bucket (0 = one launch for the whole model)")
    p.add_argument("--overlap", action="store_true", help="launch each bucket's exchange from backward on a side stream")
    p.add_argument("--error-feedback", action="store_true")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--e2e-lag", type=int, default=0, choices=[0, 1],
                   help="end-to-end loop: 1 = read step i-1's loss while step i runs (every step's loss is still read inside the "
                        "timed region; experimental, default 0 = blocking read of step i's loss)")
    p.add_argument("--no-flush", action="store_true")
    p.add_argument("--no-graph", action="store_true", help="do not capture the step in a CUDA graph")
    return p.parse_args()


def _load_clock_sampler():
    # bench/clocks.py by path: `import bench` would resolve to this very file, and the reference arm must not import ewdml
    import importlib.util
    spec = importlib.util.spec_from_file_location("ewdml_bench_clocks", os.path.join(ROOT, "bench", "clocks.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.ClockSampler


ClockSampler = _load_clock_sampler()       # NVML poller in a child process, ~2 ms period, started before warm-up


def metric_name(a) -> str:
    """One metric string for every arm (ewdml / reference / nccl); what differs between arms lives in ``config``."""
    return f"{a.model} images/sec"


def run_ewdml(a):
    import torch
    import torch.distributed as dist
    from ewdml.config import TrainConfig
    from ewdml.data import DeviceFeeder, SyntheticImageDataset
    from ewdml.parallel.peer import PeerGroup
    from ewdml.runtime import lib as L
    from ewdml.train.gpu import GpuTrainer

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != a.gpus:
        if world == 1 and a.gpus > 1:
            raise SystemExit("launch with: python -m torch.distributed.run --nproc-per-node N bench.py --gpus N ...")
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)

    dataset = "MNIST" if a.model.lower() in ("lenet", "mnistnet") else "Cifar10"
    cfg = TrainConfig(network=a.model, dataset=dataset, batch_size=a.batch_size, lr=0.01, momentum=0.9,
                      compressor=a.compressor, topk_ratio=a.topk_ratio, qsgd_levels=a.qsgd_levels,
                      exchange=a.exchange, dtype=a.dtype, compute=a.compute, enable_gpu=True, seed=1,
                      bucket_size=a.bucket_size, overlap=a.overlap, error_feedback=a.error_feedback).validate()
    pg = PeerGroup.from_env(device)
    role = "server" if (a.exchange == "ps" and rank == 0) else "worker"
    tr = GpuTrainer(cfg, device, pg, role=role)
    n_workers = world - 1 if a.exchange == "ps" else world

    ds = SyntheticImageDataset(dataset, 4096, train=True, seed=1234)
    feeder = DeviceFeeder(ds.data, ds.targets, a.batch_size, device, dtype=torch.float32, channels_last=True,
                          seed=1, shard_rank=rank, num_shards=world)
    flush = None if a.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device=device)
    x0, y0 = feeder.next()
    x0, y0 = x0.clone(), y0.clone()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_step(x, y):
        if flush is not None:
            flush.zero_()
        if role == "server":
            tr.server_step()
            return None
        loss, _ = tr.step(x, y)
        return loss

    def timed(nsteps, e2e):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        host_loss = torch.zeros(1, dtype=torch.float32).pin_memory()
        lag_buf = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(2)]
        lag_ev = [torch.cuda.Event() for _ in range(2)]
        barrier()
        t0 = time.perf_counter()
        ev0.record()
        last = None
        for _ in range(nsteps):
            if e2e:
                x, y = feeder.next()
                loss = one_step(x, y)
                feeder.mark_consumed()
                if loss is not None and not a.e2e_lag:
                    host_loss.copy_(loss.reshape(1), non_blocking=False)   # D2H read of the step's result
                    last = float(host_loss[0])
                elif loss is not None:
                    k = _ & 1                                              # enqueue this step's read, consume the previous step's
                    lag_buf[k].copy_(loss.reshape(1), non_blocking=True)
                    lag_ev[k].record()
                    if _ > 0:
                        lag_ev[k ^ 1].synchronize()
                        last = float(lag_buf[k ^ 1][0])
            else:
                one_step(x0, y0)
        if e2e and a.e2e_lag and nsteps > 0 and role == "worker":
            lag_ev[(nsteps - 1) & 1].synchronize()                         # the last step's loss, still inside the timed region
            last = float(lag_buf[(nsteps - 1) & 1][0])
        ev1.record()
        barrier()
        wall = (time.perf_counter() - t0) * 1e3
        ms = ev0.elapsed_time(ev1)
        t = torch.tensor([ms, wall], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t[1]), last

    # ---- warm-up (eager), then optional CUDA-graph capture of the whole step, then graph warm-up -----
    sampler = ClockSampler(local_rank).start() if rank == 0 else None      # running from before warm-up on
    pg.host_barrier("bench_start")          # every rank has built its trainer: nobody spins through a peer's set-up
    for _ in range(max(3, a.warmup)):
        one_step(x0, y0)
    barrier()
    if not a.no_graph:
        if role == "worker":
            tr.capture(x0, y0)              # runs 3 eager steps on a side stream, then records the graph
        else:
            for _ in range(3):              # the parameter server takes part in those 3 exchanges
                one_step(x0, y0)
        for _ in range(max(3, a.warmup)):
            one_step(x0, y0)
        barrier()
    err = tr.check_device_error()
    if err:
        raise SystemExit(f"device error word {err:#x} after warm-up")

    l0 = tr.own_launches()
    tw0 = time.time()
    ms, wall_ms, _ = timed(a.steps, e2e=False)
    tw1 = time.time()
    launches = tr.own_launches() - l0
    if sampler:
        sampler.window("device_timed", tw0, tw1)
    stats = tr.exchange.step  # noqa: F841 (keep reference alive)
    e2e = None
    if not a.no_e2e:
        for _ in range(3):
            x, y = feeder.next(); one_step(x, y); feeder.mark_consumed()
        te0 = time.time()
        ms_e, wall_e, last_loss = timed(a.steps, e2e=True)
        if sampler:
            sampler.window("e2e", te0, time.time())
        e2e = {"value": a.batch_size * n_workers * a.steps / (wall_e / 1e3), "unit": "images/s",
               "h2d_bytes_per_step": feeder.h2d_bytes_per_step, "d2h_bytes_per_step": 4,
               "ms_per_step": wall_e / a.steps, "timed_with": "host wall clock around barrier+synchronize, max over ranks", "loss_read_lag_steps": a.e2e_lag,
               "last_loss": last_loss}
    # exchange statistics (one extra step with stats collection, outside the timed regions)
    st = None
    for _ in range(3):          # three eager steps, the last one is reported: the first eager launch after graph replays pays one-off costs
        if role == "worker":
            _, st = tr.step(x0, y0, want_stats=True)
        else:
            st = tr.server_step(want_stats=True)
    barrier()
    clocks = sampler.stop() if sampler else None
    # selection diagnostics of that last exchange: segments whose sampled band missed the k-th magnitude (exact-scan fallback)
    sel = None
    try:
        plans = getattr(tr.exchange, "plans", None) or [getattr(tr.exchange, "push", None)]
        infos = [i for p in plans if p is not None for i in p.seginfo()]
        fb = torch.tensor([float(sum(p.fallbacks() for p in plans if p is not None))], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(fb)
        sel = {"segments": len(infos), "fallback_segments": sum(i["fallback"] for i in infos),
               "fallbacks_all_ranks_whole_run": int(fb.item()), "exchanges_whole_run": tr.step_idx,
               "tie_segments": sum(1 for i in infos if i["need_eq"]), "max_candidates": max((i["ncand"] for i in infos), default=0)}
    except Exception:  # noqa: BLE001 -- diagnostics only
        pass
    err = tr.check_device_error()
    value = a.batch_size * n_workers * a.steps / (ms / 1e3)
    out = {
        "metric": metric_name(a),
        "value": value, "unit": "images/s", "n_gpus": a.gpus, "steps": a.steps, "warmup": max(3, a.warmup),
        "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None,      # BASELINE.md publishes no images/s figure (CPU-only Colab minutes and MB/iteration)
        "dtype": a.dtype, "data": "synthetic (%s shape, random-init weights)" % ("MNIST" if a.model.lower() in ("lenet", "mnistnet") else "CIFAR-10"), "impl": "ewdml",
        "config": {"model": a.model, "global_batch": a.batch_size * n_workers, "per_gpu_batch": a.batch_size,
                   "seq_len": None, "image": list(tr.input_spec()[1:]),
                   "parallelism": f"dp{world}" + ("(1 server)" if a.exchange == "ps" else ""),
                   "compressor": a.compressor, "topk_ratio": a.topk_ratio, "qsgd_levels": a.qsgd_levels,
                   "exchange": a.exchange, "compute": a.compute, "cuda_graph": tr.graph is not None,
                   "bucket_size": a.bucket_size, "overlap": tr.sched is not None, "error_feedback": a.error_feedback,
                   "l2": "none" if a.no_flush else "256 MiB buffer overwritten every step inside the timed region"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "device_ms_total": ms, "wall_ms_total": wall_ms,
        "exchange": None if st is None else {"bytes_sent_per_step": st.bytes_sent, "bytes_recv_per_step": st.bytes_recv,
                                             "dense_bytes": tr.flat.dense_nbytes(),
                                             "ms_send": st.t_send * 1e3, "ms_recv_apply": st.t_recv * 1e3, "selection": sel},
        "device_error": err,
    }
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    pg.close()


def main():
    a = parse()
    if a.impl == "reference":
        from baseline.run_reference import main as ref_main
        return ref_main(a, ClockSampler, metric_name)
    if a.impl == "nccl":
        from baseline.nccl_same_algorithm import main as nccl_main
        return nccl_main(a, ClockSampler, metric_name)
    return run_ewdml(a)


if __name__ == "__main__":
    main()
