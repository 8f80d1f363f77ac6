"""``distributed_nn.py`` entry point -- same flags and env contract as the reference.

``python -m torch.distributed.run --nproc-per-node=3 distributed_nn.py --network=LeNet
--dataset=MNIST ...`` (``src/run_pytorch_single.sh:1-17``).  Reads ``RANK/WORLD_SIZE/MASTER_ADDR/
MASTER_PORT`` (``distributed_nn.py:75-78``); rank 0 is the parameter server and ranks 1..N-1 are
workers when ``--exchange ps`` (the default and the reference's only mode); with
``--exchange allgather|allreduce`` every rank is a worker.

Backend selection (``--backend``): ``gloo`` = CPU plumbing (BASELINE.json config 1); ``native`` =
the hand-written NVLink kernels over the symmetric peer heap (needs one B200 per rank); ``nccl`` =
the same algorithm on stock NCCL collectives (the measured baseline); ``auto`` picks ``native`` when
``--enable-gpu`` and a GPU per rank is visible, else ``gloo``.
"""
from __future__ import annotations

import os
import time

import torch
import torch.distributed as dist

from ..config import parse_args
from ..data import build_datasets, prepare_data
from ..ops import layers as L
from ..train.roles import DistributedWorker, SyncReplicasMaster_NN, planned_steps
from ..utils.metrics import logger


def _pick_backend(cfg, world_size: int) -> str:
    if cfg.backend != "auto":
        return cfg.backend
    if cfg.enable_gpu and not cfg.no_cuda and torch.cuda.is_available() \
            and torch.cuda.device_count() >= min(world_size, 8):
        return "native"
    return "gloo"


def main(argv=None):
    cfg = parse_args(argv, "ewdml parameter-server / data-parallel trainer")
    rank = int(os.environ.get("RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", cfg.local_rank if cfg.local_rank is not None else rank))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    print(rank, world_size)

    backend = _pick_backend(cfg, world_size)
    use_gpu = backend in ("native", "nccl") or (cfg.enable_gpu and torch.cuda.is_available())
    device = torch.device("cuda", local_rank % max(1, torch.cuda.device_count())) if use_gpu else torch.device("cpu")
    if use_gpu:
        torch.cuda.set_device(device)
    torch.manual_seed(cfg.seed)
    L.set_compute(cfg.compute if use_gpu else "torch")

    dist.init_process_group(backend="nccl" if backend == "nccl" else "gloo",
                            world_size=world_size, rank=rank)

    exchange_factory = None
    if backend == "native":
        from ..parallel.native_exchange import make_native_factory
        exchange_factory = make_native_factory(device)
    elif backend == "nccl":
        from ..parallel.nccl_baseline import build_nccl_exchange as exchange_factory

    ps_mode = cfg.exchange == "ps"
    if ps_mode and world_size < 2:
        raise SystemExit("--exchange ps needs WORLD_SIZE >= 2 (rank 0 is the server)")
    n_workers = world_size - 1 if ps_mode else world_size
    if backend != "gloo" and rank == 0 and (cfg.mode != "normal" or (ps_mode and 0 < cfg.num_aggregate < n_workers)):
        # backup workers, the straggler cut-off and the asynchronous server live in the gloo parameter server only
        logger.warning("--mode %s / --num-aggregate %d act on --backend gloo only: the %s backend waits for all %d workers "
                       "every step", cfg.mode, cfg.num_aggregate, backend, n_workers)
    shard = (rank - 1) if ps_mode else rank
    train_set, _ = build_datasets(cfg)
    total = planned_steps(cfg, len(train_set), n_workers)

    t0 = time.time()
    if backend == "native" and cfg.graphs:
        # fast path: GpuTrainer + DeviceFeeder, the whole step replayed from one CUDA graph
        from ..train.gpu import run_training
        res = run_training(cfg, rank, world_size, device, total_steps=total)
        if ps_mode and rank == 0:
            print("Done sending messages to workers! , total time takken by master " + str(time.time() - t0))
        else:
            print("Now the next step is: {}".format(total))
            print("totla time by worker: {} is {} ".format(rank, time.time() - t0))
        dist.barrier()
        dist.destroy_process_group()
        return res
    if ps_mode and rank == 0:
        master = SyncReplicasMaster_NN(cfg, world_size, device, exchange_factory=exchange_factory)
        master.build_model(num_classes=cfg.num_classes())
        print("I am the master: the world size is {}, cur step: {}".format(world_size, master.cur_step))
        master.start(total_steps=total)
        print("Done sending messages to workers! , total time takken by master " + str(time.time() - t0))
    else:
        train_loader, test_loader = prepare_data(cfg, shard_rank=shard, num_shards=n_workers)
        worker = DistributedWorker(cfg, rank, world_size, device, exchange_factory=exchange_factory)
        worker.build_model(num_classes=cfg.num_classes())
        print("I am worker: {} in all {} workers, next step: {}".format(rank, n_workers, worker.next_step))
        worker.train(train_loader, test_loader, total_steps=total)
        print("Now the next step is: {}".format(worker.next_step))
        print("totla time by worker: {} is {} ".format(rank, time.time() - t0))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
