"""Training roles: worker, parameter server ("master") and the shared single-process base.

Counterparts: ``DistributedWorker`` (``src/distributed_worker.py:62-398``),
``SyncReplicasMaster_NN`` (``src/sync_replicas_master_nn.py:89-266``) and ``NN_Trainer``
(``src/nn_ops.py:28-105``).  The per-step structure is that of ``train_updated`` /
``start_updated``; the exchange itself is delegated to an ``Exchange``
strategy (host collectives or the native NVLink kernels), selected by ``--exchange/--backend``.

Fixed relative to the reference: initial weights are broadcast from rank 0, workers read disjoint
shards, seeding is applied, one rank writes checkpoints atomically, and all ranks agree on the
number of steps so nobody dead-locks on a collective when an epoch ends.
"""
from __future__ import annotations

import os
import time
from typing import Optional

import torch
import torch.distributed as dist

from ..config import TrainConfig
from ..models import build_model
from ..ops import layers as L
from ..optim import build_optimizer
from ..parallel.flat import FlatModel
from ..parallel.host_exchange import build_host_exchange
from ..utils import checkpoint as ckpt
from ..utils.metrics import StepLogger, accuracy, logger


def planned_steps(cfg: TrainConfig, dataset_len: int, num_shards: int) -> int:
    per_epoch = max(1, (dataset_len // max(1, num_shards if cfg.shard_data else 1)) // cfg.batch_size)
    return min(cfg.max_steps, cfg.epochs * per_epoch)


class NNTrainer:
    """Single-process trainer (``NN_Trainer.train_and_validate``, ``nn_ops.py:47-105``)."""

    def __init__(self, cfg: TrainConfig, device: Optional[torch.device] = None):
        self.cfg = cfg
        self.device = device or torch.device("cpu")
        self.cur_step = 0

    def build_model(self, num_classes: int = 10):
        torch.manual_seed(self.cfg.seed)
        self.network = build_model(self.cfg.network, num_classes)
        self.flat = FlatModel(self.network, self.device,
                              channels_last=self.device.type == "cuda")
        self.optimizer = build_optimizer(self.cfg, self.flat.params)
        self.criterion = L.cross_entropy
        if self.cfg.resume:
            self._resume(self.cfg.resume)
        return self

    # -- helpers shared by the roles ---------------------------------------------------
    def _forward(self, x, y):
        logits = self.network(x)
        if getattr(self.network, "log_softmax_output", False):
            return torch.nn.functional.nll_loss(logits, y), logits
        return self.criterion(logits, y), logits

    def _train_init(self):
        self.network.train()
        self.flat.zero_grad()

    def _resume(self, path: str):
        ckpt.load_model(self.network, path)
        side = ckpt.load_sidecar(path)
        if side is not None:
            self.cur_step = int(side["step"])
            if hasattr(self, "next_step"):
                self.next_step = self.cur_step          # what the worker announces before it runs its first step
            if side.get("optimizer"):
                try:
                    self.optimizer.load_state_dict(side["optimizer"])
                except Exception as e:  # optimizer shape changed: keep weights, restart momentum
                    logger.warning("could not restore optimizer state: %s", e)
            ex = side.get("extra") or {}
            if ex.get("momentum") is not None and getattr(self.flat, "momentum", None) is not None:
                self.flat.momentum.copy_(ex["momentum"])
            exch = getattr(self, "exchange", None)
            if ex.get("residual") is not None and getattr(exch, "residual", None) is not None:
                exch.residual.copy_(ex["residual"])
            if side.get("torch_rng") is not None:
                # only the rank that wrote the sidecar continues its own generator; the others re-derive theirs
                if side.get("rng_rank") in (None, getattr(self, "rank", None)):
                    torch.set_rng_state(side["torch_rng"])
                else:
                    ckpt.reseed_for_rank(self.cfg.seed, getattr(self, "rank", 0), self.cur_step)
        logger.info("resumed from %s at step %d", path, self.cur_step)

    def _generate_model_path(self) -> str:
        return ckpt.model_path(self.cfg.train_dir, self.cur_step, self.cfg.legacy_checkpoint_name)

    def _save_model(self, file_path: str, is_writer: bool = True):
        crc = ckpt.save_model(self.network, file_path, is_writer)
        exch = getattr(self, "exchange", None)
        mom = getattr(self.flat, "momentum", None)
        res = getattr(exch, "residual", None)
        extra = {"momentum": mom.detach().cpu() if mom is not None else None,           # flat-buffer optimizers keep their
                 "residual": res.detach().cpu() if res is not None else None}          # state outside optimizer.state_dict()
        ckpt.save_sidecar(file_path, step=self.cur_step, optimizer_state=self.optimizer.state_dict(), extra=extra,
                          is_writer=is_writer, model_crc=crc, rng_rank=getattr(self, "rank", None))

    @torch.no_grad()
    def _evaluate_model(self, test_loader):
        self.network.eval()
        loss_sum, n, p1, p5, batches = 0.0, 0, 0.0, 0.0, 0
        for x, y in test_loader:
            x, y = x.to(self.device), y.to(self.device)
            out = self.network(x).float()
            lp = out if getattr(self.network, "log_softmax_output", False) else torch.log_softmax(out, 1)
            loss_sum += torch.nn.functional.nll_loss(lp, y, reduction="sum").item()
            a1, a5 = accuracy(out, y, topk=(1, 5))
            p1 += a1.item(); p5 += a5.item(); batches += 1; n += y.shape[0]
        res = {"step": self.cur_step, "loss": loss_sum / max(1, n), "prec1": p1 / max(1, batches),
               "prec5": p5 / max(1, batches)}
        print("Test set: Step: {step}, Average loss: {loss:.4f}, Prec@1: {prec1} Prec@5: {prec5}".format(**res))
        return res

    def train_and_validate(self, train_loader, test_loader):
        for _ in range(self.cfg.epochs):
            for x, y in train_loader:
                if self.cur_step >= self.cfg.max_steps:
                    return
                self._train_init()
                loss, _ = self._forward(x.to(self.device), y.to(self.device))
                loss.backward()
                self.optimizer.step(grads=self.flat.grad_views())
                self.cur_step += 1
            self._evaluate_model(test_loader)


class DistributedWorker(NNTrainer):
    def __init__(self, cfg: TrainConfig, rank: int, world_size: int,
                 device: Optional[torch.device] = None, group=None, exchange_factory=None):
        super().__init__(cfg, device)
        self.rank, self.world_size, self.group = rank, world_size, group
        self.next_step = 0
        self._exchange_factory = exchange_factory or build_host_exchange
        self.log = StepLogger(rank, cfg.metrics_file, cfg.log_interval)
        self.history = []

    def build_model(self, num_classes: int = 10):
        super().build_model(num_classes)
        self.exchange = self._exchange_factory("worker", self.flat, self.cfg, self.optimizer, self.group)
        self.exchange.broadcast_initial_weights(0)
        # --overlap: fire each bucket's exchange from the backward pass (LeNetSplit's per-layer Isend, lenet.py:111-186,
        # generalised to every model); only exchanges that expose the overlapped form take part
        self.sched = None
        if self.cfg.overlap and hasattr(self.exchange, "begin_overlap") and hasattr(self.exchange, "buckets") \
                and self.cfg.resolved_sync_period() == 1 and not hasattr(self.exchange, "state"):
            from ..parallel.overlap import BucketScheduler
            self.sched = BucketScheduler(self.flat.params, self.exchange.buckets, self.exchange.bucket_ready).install()
        return self

    @property
    def is_checkpoint_writer(self) -> bool:
        first_worker = 1 if self.cfg.exchange == "ps" else 0
        return self.rank == first_worker

    def train(self, train_loader, test_loader=None, total_steps: Optional[int] = None):
        cfg = self.cfg
        period = cfg.resolved_sync_period()
        total = total_steps if total_steps is not None else cfg.max_steps
        logger.info("Worker %d: starting training", self.rank)
        done = False
        for epoch in range(cfg.epochs):
            if hasattr(getattr(train_loader, "sampler", None), "set_epoch"):
                train_loader.sampler.set_epoch(epoch)
            for x, y in train_loader:
                if self.cur_step >= total:
                    done = True
                    break
                if (cfg.fault_inject_step == self.cur_step and cfg.fault_inject_rank == self.rank
                        and os.environ.get("EWDML_RESTART_COUNT", "0") == "0"):       # once: not again after a supervised restart
                    logger.error("fault injection: rank %d exits at step %d", self.rank, self.cur_step)
                    os._exit(17)
                if cfg.straggler_rank == self.rank and cfg.straggler_delay_s > 0:
                    time.sleep(cfg.straggler_delay_s)
                self._train_init()
                t0 = time.perf_counter()
                x, y = x.to(self.device), y.to(self.device)
                loss, logits = self._forward(x, y)
                if self.sched is not None:
                    self.exchange.begin_overlap(self.cur_step)
                    self.sched.begin_step()
                    loss.backward()
                    self.sched.flush()
                    t_comp = time.perf_counter() - t0
                    st = self.exchange.finish_overlap()
                else:
                    loss.backward()
                    t_comp = time.perf_counter() - t0
                    sync = self.cur_step % period == 0
                    if hasattr(self.exchange, "state"):     # native kernels: payload size / phase times cost a device sync,
                        st = self.exchange.step(self.cur_step, is_sync_step=sync,      # so only on the steps that get logged
                                                want_stats=self.cur_step % max(1, cfg.log_interval) == 0)
                    else:
                        st = self.exchange.step(self.cur_step, is_sync_step=sync)
                p1, p5 = accuracy(logits.detach().float(), y, topk=(1, 5))
                rec = self.log.log(self.cur_step, loss=float(loss.detach()), prec1=float(p1), prec5=float(p5),
                                   bytes_sent=st.bytes_sent, bytes_recv=st.bytes_recv,
                                   dense_bytes=self.flat.dense_nbytes(), t_send=st.t_send,
                                   t_recv=st.t_recv, t_compute=t_comp, images=int(y.shape[0]))
                self.history.append(rec)
                finished = self.cur_step
                self.cur_step += 1                      # from here on: the NEXT step to run (what the sidecar records)
                self.next_step = self.cur_step
                if finished % cfg.eval_freq == 0:
                    self._save_model(ckpt.model_path(cfg.train_dir, finished, cfg.legacy_checkpoint_name),
                                     self.is_checkpoint_writer)
            if done:
                break
        self.log.close()
        return self.history

    # reference spelling
    train_updated = train


class SyncReplicasMaster_NN(NNTrainer):
    def __init__(self, cfg: TrainConfig, world_size: int, device: Optional[torch.device] = None,
                 group=None, exchange_factory=None):
        super().__init__(cfg, device)
        self.world_size, self.group = world_size, group
        self._num_workers = world_size - 1
        self._exchange_factory = exchange_factory or build_host_exchange
        self.cur_step = 0

    def build_model(self, num_classes: int = 10):
        super().build_model(num_classes)
        self.exchange = self._exchange_factory("server", self.flat, self.cfg, self.optimizer, self.group)
        self.exchange.broadcast_initial_weights(0)
        return self

    def start(self, total_steps: Optional[int] = None):
        cfg = self.cfg
        period = cfg.resolved_sync_period()
        total = total_steps if total_steps is not None else cfg.max_steps
        stats = []
        if cfg.mode == "async":
            # one exchange.step per arriving gradient; every worker pushes once per step of its own
            n = (total - self.cur_step) * self._num_workers
            self.exchange.expect(total - self.cur_step)
            for i in range(n):
                stats.append(self.exchange.step(i))
            self.cur_step = total
            self.exchange.drain()
            return stats
        if hasattr(self.exchange, "expect"):
            self.exchange.expect(sum(1 for i in range(self.cur_step, total) if i % period == 0))
        for i in range(self.cur_step, total):
            if i % max(1, cfg.log_interval) == 0:
                logger.info("Master node is entering step: %d", i)
            st = self.exchange.step(i, is_sync_step=(i % period == 0))
            stats.append(st)
            self.cur_step = i + 1
            if cfg.resolved_pull() == "weights" and i % cfg.eval_freq == 0:
                # the server owns the authoritative weights in this mode
                ckpt.save_model(self.network, ckpt.model_path(cfg.train_dir, i, False), True)
        if hasattr(self.exchange, "drain"):
            self.exchange.drain()
        return stats

    start_updated = start
