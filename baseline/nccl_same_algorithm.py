"""``bench.py --impl nccl``: the same algorithm on stock PyTorch + NCCL (the measured baseline).

BASELINE.json: "a path that only calls NCCL or MPI for the gradient exchange is the baseline, not the
product".  This arm is that path, made as strong as a library-only implementation reasonably gets:
PyTorch-eager model (channels-last, bf16 autocast, same as the ewdml torch-compute arm), the
reference's compressor math as torch ops per parameter tensor (``torch.topk`` on |g|,
``Compresssor/TopK.py:5-10``; QSGD levels with ``torch.rand``, ``Compresssor/qsgd.py:17-28``), ONE
coalesced ``all_gather`` per step for (levels, indices, norms), decompress + average with
``index_add_``, then ``torch.optim.SGD``.  Same metric, model, batch, steps, timing rules and JSON
line as the ewdml arm; none of this repo's kernels run here.
"""
from __future__ import annotations

import json
import os
import time

import torch
import torch.distributed as dist
import torch.nn.functional as F


def main(a):
    from ewdml.data import DeviceFeeder, SyntheticImageDataset
    from ewdml.models import build_model
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
    torch.manual_seed(1)
    dataset = "MNIST" if a.model.lower() in ("lenet", "mnistnet") else "Cifar10"
    model = build_model(a.model, 10).to(device).to(memory_format=torch.channels_last)
    opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9)
    params = [p for p in model.parameters()]
    ks = [max(1, int(p.numel() * a.topk_ratio)) for p in params]
    offs = [0]
    for k in ks:
        offs.append(offs[-1] + k)
    K = offs[-1]
    s = float(a.qsgd_levels)
    lv = torch.empty(K, dtype=torch.int16, device=device)
    ix = torch.empty(K, dtype=torch.int32, device=device)
    nm = torch.empty(len(params), dtype=torch.float32, device=device)
    g_lv = torch.empty(world * K, dtype=torch.int16, device=device)
    g_ix = torch.empty(world * K, dtype=torch.int32, device=device)
    g_nm = torch.empty(world * len(params), dtype=torch.float32, device=device)
    compress = a.compressor != "none"

    def exchange():
        if not compress:
            if world > 1:
                for p in params:
                    dist.all_reduce(p.grad)
                    p.grad /= world
            return
        for i, p in enumerate(params):
            g = p.grad.reshape(-1)
            _, idx = torch.topk(g.abs(), ks[i], sorted=False)
            v = g[idx]
            norm = v.norm()
            lf = s / norm * v.abs()
            prev = lf.floor()
            level = prev + (torch.rand_like(v) < (lf - prev))
            lv[offs[i]:offs[i + 1]] = (level * v.sign()).to(torch.int16)
            ix[offs[i]:offs[i + 1]] = idx.to(torch.int32)
            nm[i] = norm
        if world > 1:
            # NCCL has no int16: ship the level buffer as raw bytes
            dist.all_gather_into_tensor(g_lv.view(torch.uint8), lv.view(torch.uint8))
            dist.all_gather_into_tensor(g_ix, ix); dist.all_gather_into_tensor(g_nm, nm)
        else:
            g_lv.copy_(lv); g_ix.copy_(ix); g_nm.copy_(nm)
        for i, p in enumerate(params):
            g = p.grad.reshape(-1) if p.grad.is_contiguous() else None
            dense = torch.zeros(p.numel(), dtype=torch.float32, device=device)
            for r in range(world):
                q = g_lv[r * K + offs[i]: r * K + offs[i + 1]].float()
                dense.index_add_(0, g_ix[r * K + offs[i]: r * K + offs[i + 1]].long(), g_nm[r * len(params) + i] / s * q)
            p.grad.copy_((dense / world).view_as(p.grad))

    ds = SyntheticImageDataset(dataset, 4096, train=True, seed=1234)
    feeder = DeviceFeeder(ds.data, ds.targets, a.batch_size, device, dtype=torch.float32, channels_last=True, seed=1,
                          shard_rank=rank, num_shards=world)
    flush = None if a.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device=device)
    x0, y0 = feeder.next(); x0, y0 = x0.clone(), y0.clone()

    def one_step(x, y):
        if flush is not None:
            flush.zero_()
        opt.zero_grad(set_to_none=False)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=a.dtype == "bf16"):
            logits = model(x)
        loss = F.cross_entropy(logits.float(), y)
        loss.backward()
        exchange()
        opt.step()
        return loss.detach()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(n, e2e):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        host = torch.zeros(1).pin_memory()
        barrier(); t0 = time.perf_counter(); ev0.record()
        for _ in range(n):
            if e2e:
                x, y = feeder.next(); loss = one_step(x, y); feeder.mark_consumed(); host.copy_(loss.reshape(1)); _ = float(host[0])
            else:
                one_step(x0, y0)
        ev1.record(); barrier()
        t = torch.tensor([ev0.elapsed_time(ev1), (time.perf_counter() - t0) * 1e3], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t[1])

    W = max(3, a.warmup)
    for _ in range(W):
        one_step(x0, y0)
    ms, _ = timed(a.steps, False)
    for _ in range(3):
        x, y = feeder.next(); one_step(x, y); feeder.mark_consumed()
    _, wall_e = timed(a.steps, True)
    value = a.batch_size * world * a.steps / (ms / 1e3)
    if rank == 0:
        print(json.dumps({"impl": "nccl-same-algorithm", "metric": f"{a.model} images/sec ({a.compressor} k={a.topk_ratio} s={a.qsgd_levels} torch ops + NCCL all_gather)",
                          "value": value, "unit": "images/s", "n_gpus": world, "steps": a.steps, "warmup": W, "ms_per_step": ms / a.steps,
                          "higher_is_better": True, "scaling": "weak", "vs_baseline": value / (104.0 * world), "dtype": a.dtype,
                          "data": "synthetic (CIFAR-10 shape, random-init weights)",
                          "config": {"model": a.model, "global_batch": a.batch_size * world, "per_gpu_batch": a.batch_size, "parallelism": f"dp{world}",
                                     "compressor": a.compressor, "exchange": "nccl all_gather (coalesced)"},
                          "e2e": {"value": a.batch_size * world * a.steps / (wall_e / 1e3), "unit": "images/s",
                                  "h2d_bytes_per_step": feeder.h2d_bytes_per_step, "d2h_bytes_per_step": 4},
                          "bytes_sent_per_step": (world - 1) * (K * 6 + 4 * len(params)) if compress else 2 * (world - 1) / world * 4 * sum(p.numel() for p in params),
                          "gpu_launches": 0}), flush=True)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()
