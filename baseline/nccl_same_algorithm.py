"""``bench.py --impl nccl``: the strong library baseline -- stock PyTorch/cuDNN + NCCL, CUDA-graphed.

BASELINE.json: "a path that only calls NCCL or MPI for the gradient exchange is the baseline, not the
product".  The reference's own GPU build (Horovod + Open MPI + NCCL, ``horvod_pytorch.py:173-201``) cannot be
built on this box, so this arm is what that build does, made as strong as a library-only implementation gets:

* model: the same architecture, channels-last, bf16 autocast over fp32 master weights -> cuDNN / cuBLAS kernels;
* gradients accumulate straight into ONE flat fp32 buffer (every ``p.grad`` is a view of it);
* ``--compressor none``: Horovod-style fused allreduce -- ``dist.all_reduce`` over 32 MB slices of the flat buffer
  (``Horovod all reduce.ipynb`` cell 10, ``--fusion-threshold-mb 32``), 1/W folded into the optimizer step;
* ``--compressor topk-qsgd | topk | qsgd``: the reference's compressor math (``Compresssor/TopK.py:5-10``,
  ``Compresssor/qsgd.py:17-28``) with per-tensor ``torch.topk`` and per-tensor norms, everything after the select done
  on flat buffers, ONE ``all_gather_into_tensor`` of (int16 levels | int32 indices | fp32 norms), one ``index_add_``
  to decompress + average;
* flat momentum-SGD with ``torch._foreach`` ops;
* the WHOLE step (forward, backward, exchange, optimizer) captured in one CUDA graph (``--no-graph`` = eager).

Same metric string, model, batch, step counts, timing rules, clocks record and ``e2e`` region (pinned-host ->
device copy of every batch, device -> host read of every loss) as the ewdml arm.  None of this repo's kernels run
here (``ewdml.models`` layers execute their plain torch path; libewdml.so is not loaded).
"""
from __future__ import annotations

import json
import os
import time

import torch
import torch.distributed as dist
import torch.nn.functional as F

FUSION_BYTES = 32 << 20


def main(a, ClockSampler, metric_name):
    from ewdml.data import DeviceFeeder, SyntheticImageDataset
    from ewdml.models import build_model
    from ewdml.ops import layers as Lyr
    Lyr.set_compute("torch")
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
    torch.manual_seed(1)
    torch.backends.cudnn.benchmark = True
    # --dtype: bf16 = autocast over fp32 master weights; tf32 = fp32 tensors, tensor-core convolutions and GEMMs with
    # 10-bit mantissa inputs (the fastest library mode that keeps the reference's fp32 storage); fp32 = no tensor cores.
    if a.dtype not in ("bf16", "tf32", "fp32"):
        raise SystemExit(f"--impl nccl: --dtype must be bf16, tf32 or fp32, got {a.dtype}")
    torch.backends.cuda.matmul.allow_tf32 = a.dtype == "tf32"
    torch.backends.cudnn.allow_tf32 = a.dtype == "tf32"
    dataset = "MNIST" if a.model.lower() in ("lenet", "mnistnet") else "Cifar10"
    model = build_model(a.model, 10).to(device).to(memory_format=torch.channels_last)
    model.train()
    params = [p for p in model.parameters()]
    numels = [p.numel() for p in params]
    P = sum(numels)
    flat_g = torch.zeros(P, dtype=torch.float32, device=device)
    o = 0
    poff = []
    for p in params:
        p.grad = flat_g[o:o + p.numel()].view(p.shape)      # autograd accumulates in place into the flat buffer
        poff.append(o); o += p.numel()
    grads = [p.grad for p in params]
    mom = [torch.zeros_like(p) for p in params]
    lr, mu = 0.01, 0.9
    compress = a.compressor != "none"
    use_topk = a.compressor in ("topk", "topk-qsgd")
    use_qsgd = a.compressor in ("qsgd", "topk-qsgd")
    s = float(a.qsgd_levels)

    # ---- flat payload layout: [int16 levels or fp32 values | int32 indices | fp32 norms] --------------------------
    ks = [max(1, int(n * a.topk_ratio)) if use_topk else n for n in numels]
    koff = [0]
    for k in ks:
        koff.append(koff[-1] + k)
    K = koff[-1]
    nseg = len(params)
    val_bytes = 2 if use_qsgd else 4
    b_val = (K * val_bytes + 15) // 16 * 16
    b_idx = (K * 4 + 15) // 16 * 16 if use_topk else 0
    b_nrm = (nseg * 4 + 15) // 16 * 16
    nbytes = b_val + b_idx + b_nrm
    if compress:
        pay = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        allpay = torch.zeros(world, nbytes, dtype=torch.uint8, device=device)
        vals = torch.zeros(K, dtype=torch.float32, device=device)
        idx64 = torch.zeros(K, dtype=torch.int64, device=device)
        seg_of = torch.repeat_interleave(torch.arange(nseg, device=device), torch.tensor(ks, device=device))
        ent_base = torch.repeat_interleave(torch.tensor(poff, device=device), torch.tensor(ks, device=device))
        dense = torch.zeros(P, dtype=torch.float32, device=device)
        sq = torch.zeros(nseg, dtype=torch.float32, device=device)

    def exchange():
        if world == 1 and not compress:
            return
        if not compress:
            for lo in range(0, P, FUSION_BYTES // 4):
                dist.all_reduce(flat_g[lo:lo + FUSION_BYTES // 4])
            return                                           # the 1/W lives in the optimizer step
        if use_topk:
            for i, g in enumerate(grads):
                gf = g.reshape(-1)
                torch.topk(gf.abs(), ks[i], sorted=False, out=(vals[koff[i]:koff[i + 1]], idx64[koff[i]:koff[i + 1]]))
                torch.index_select(gf, 0, idx64[koff[i]:koff[i + 1]], out=vals[koff[i]:koff[i + 1]])
            v = vals
            gidx = idx64 + ent_base
            pay[b_val:b_val + K * 4].view(torch.int32).copy_(gidx)
        else:
            v = flat_g
        norms = pay[b_val + b_idx:b_val + b_idx + nseg * 4].view(torch.float32)
        if use_qsgd:
            sq.zero_().index_add_(0, seg_of, v * v)
            torch.sqrt(sq, out=norms)
            lf = v.abs() * (s / norms.clamp_min(1e-30))[seg_of]
            prev = lf.floor()
            level = (prev + (torch.rand_like(lf) < (lf - prev))).clamp_(max=s)
            pay[:K * 2].view(torch.int16).copy_(level * v.sign())
        else:
            norms.fill_(1.0)
            pay[:K * 4].view(torch.float32).copy_(v)
        if world > 1:
            dist.all_gather_into_tensor(allpay.view(-1), pay)
        else:
            allpay[0].copy_(pay)
        # decompress every sender with its own norm, average, back into the flat gradient
        if use_qsgd:
            q = allpay[:, :K * 2].contiguous().view(torch.int16).view(world, K).float()
            nr = allpay[:, b_val + b_idx:b_val + b_idx + nseg * 4].contiguous().view(torch.float32).view(world, nseg)
            vv = q * (nr / s)[:, seg_of]
        else:
            vv = allpay[:, :K * 4].contiguous().view(torch.float32).view(world, K)
        if use_topk:
            gi = allpay[:, b_val:b_val + K * 4].contiguous().view(torch.int32).view(world, K).long()
            dense.zero_().index_add_(0, gi.reshape(-1), vv.reshape(-1))
            torch.mul(dense, 1.0 / world, out=flat_g)
        else:
            torch.mul(vv.sum(0), 1.0 / world, out=flat_g)

    gscale = (1.0 / world) if not compress else 1.0

    def sgd():
        torch._foreach_mul_(mom, mu)
        torch._foreach_add_(mom, grads, alpha=gscale)
        torch._foreach_add_(params, mom, alpha=-lr)

    def body(x, y):
        flat_g.zero_()
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=a.dtype == "bf16"):
            logits = model(x)
        loss = F.cross_entropy(logits.float(), y)
        loss.backward()
        exchange()
        with torch.no_grad():
            sgd()
        return loss.detach()

    ds = SyntheticImageDataset(dataset, 4096, train=True, seed=1234)
    feeder = DeviceFeeder(ds.data, ds.targets, a.batch_size, device, dtype=torch.float32, channels_last=True, seed=1,
                          shard_rank=rank, num_shards=world)
    flush = None if a.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device=device)
    x0, y0 = feeder.next(); x0, y0 = x0.clone(), y0.clone()
    sx, sy = x0.clone(), y0.clone()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    W = max(3, a.warmup)
    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    graph, static_loss = None, None
    side = torch.cuda.Stream(device=device)
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(max(W, 11)):          # eager warm-up on a side stream (cuDNN autotune, NCCL communicator set-up)
            body(sx, sy)
    torch.cuda.current_stream().wait_stream(side)
    barrier()
    if not a.no_graph:
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            static_loss = body(sx, sy)
        barrier()

    def one_step(x, y):
        if flush is not None:
            flush.zero_()
        if graph is None:
            return body(x, y)
        sx.copy_(x, non_blocking=True); sy.copy_(y, non_blocking=True)
        graph.replay()
        return static_loss

    def timed(n, e2e):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        host = torch.zeros(1).pin_memory()
        last = None
        barrier(); tw0 = time.time(); t0 = time.perf_counter(); ev0.record()
        for _ in range(n):
            if e2e:
                x, y = feeder.next(); loss = one_step(x, y); feeder.mark_consumed()
                host.copy_(loss.reshape(1)); last = float(host[0])
            else:
                one_step(x0, y0)
        ev1.record(); barrier()
        t = torch.tensor([ev0.elapsed_time(ev1), (time.perf_counter() - t0) * 1e3], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t[1]), last, (tw0, time.time())

    for _ in range(W):
        one_step(x0, y0)
    ms, _, _, win = timed(a.steps, False)
    if sampler:
        sampler.window("device_timed", *win)
    for _ in range(3):
        x, y = feeder.next(); one_step(x, y); feeder.mark_consumed()
    _, wall_e, last_loss, win = timed(a.steps, True)
    if sampler:
        sampler.window("e2e", *win)
    clocks = sampler.stop() if sampler else None
    value = a.batch_size * world * a.steps / (ms / 1e3)
    if rank == 0:
        sent = (world - 1) * nbytes if compress else int(2 * (world - 1) / world * 4 * P)
        print(json.dumps({"impl": "nccl", "metric": metric_name(a),
                          "value": value, "unit": "images/s", "n_gpus": world, "steps": a.steps, "warmup": W, "ms_per_step": ms / a.steps,
                          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": a.dtype,
                          "data": "synthetic (%s shape, random-init weights)" % ("MNIST" if a.model.lower() in ("lenet", "mnistnet") else "CIFAR-10"),
                          "config": {"model": a.model, "global_batch": a.batch_size * world, "per_gpu_batch": a.batch_size, "parallelism": f"dp{world}",
                                     "compressor": a.compressor, "topk_ratio": a.topk_ratio, "qsgd_levels": a.qsgd_levels,
                                     "compute": {"bf16": "torch/cuDNN channels-last, bf16 autocast", "tf32": "torch/cuDNN channels-last, fp32 tensors with TF32 tensor cores",
                                                 "fp32": "torch/cuDNN channels-last fp32, TF32 off"}[a.dtype],
                                     "exchange": ("nccl all_gather_into_tensor, one call" if compress else "nccl all_reduce, 32 MB fused slices"),
                                     "cuda_graph": graph is not None,
                                     "l2": "none" if a.no_flush else "256 MiB buffer overwritten every step inside the timed region"},
                          "e2e": {"value": a.batch_size * world * a.steps / (wall_e / 1e3), "unit": "images/s",
                                  "h2d_bytes_per_step": feeder.h2d_bytes_per_step, "d2h_bytes_per_step": 4,
                                  "ms_per_step": wall_e / a.steps, "last_loss": last_loss},
                          "clocks": clocks, "bytes_sent_per_step": sent, "gpu_launches": 0}), flush=True)
    if world > 1:
        # tearing the process group down while a CUDA graph that captured NCCL kernels is alive hangs in destroy_process_group
        # (seen at 2 GPUs): drop the graph first, and leave without the collective teardown once every rank has printed
        import sys
        sys.stdout.flush()
        dist.barrier()
        graph = None
        torch.cuda.synchronize()
        os._exit(0)
