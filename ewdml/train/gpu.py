"""GPU training engine: the public API behind ``bench.py`` and ``distributed_nn.py --enable-gpu``.

One process per GPU.  A step is: wait for the prefetched batch (pinned host -> device copy issued
one step earlier on a copy stream) -> forward/backward (bf16 autocast over fp32 master weights on
the torch compute path, or the hand-written sm_100a kernels with ``compute="native"``) writing
gradients straight into the flat gradient buffer -> exchange kernels (select + QSGD + pack + P2P
stores, then unpack + average + momentum-SGD) on the same stream.  Nothing in the step synchronises
with the host; the loss is read back only when the caller asks for it.

This is the B200 equivalent of ``DistributedWorker.train_updated`` (``distributed_worker.py:162-239``)
with its per-tensor D2H copies, blocking gloo collectives and Python optimizer loop removed.
"""
from __future__ import annotations

from typing import Optional

import torch

from ..config import TrainConfig
from ..models import build_model, input_shape
from ..ops import layers as Lyr
from ..parallel.flat import FlatModel
from ..parallel.native_exchange import (NativeAllgather, NativeAllreduce, NativePSServer, NativePSWorker,
                                        native_allocator)
from ..parallel.peer import PeerGroup
from ..runtime import lib as L


class GpuTrainer:
    def __init__(self, cfg: TrainConfig, device: torch.device, pg: Optional[PeerGroup] = None,
                 role: str = "worker"):
        self.cfg, self.device, self.role = cfg, device, role
        torch.cuda.set_device(device)
        self.pg = pg or PeerGroup(0, 1, device)
        self.rank, self.world = self.pg.rank, self.pg.world
        torch.manual_seed(cfg.seed)                      # identical initial weights on every rank
        Lyr.set_compute(cfg.compute)
        self.model = build_model(cfg.network, cfg.num_classes())
        hint = {}
        self.flat = FlatModel(self.model, device, channels_last=True, with_bf16=(cfg.compute == "native"),
                              allocator=native_allocator(self.pg, hint) if cfg.exchange == "allreduce" else None)
        self.flat.grad_sym = hint.get("grad_sym")
        for b in self.model.buffers():
            b.data = b.data.to(device)
        if cfg.exchange == "ps":
            cls = NativePSServer if role == "server" else NativePSWorker
            self.exchange = cls(self.flat, cfg, None, self.pg)
        elif cfg.exchange == "allgather":
            self.exchange = NativeAllgather(self.flat, cfg, None, self.pg)
        else:
            self.exchange = NativeAllreduce(self.flat, cfg, None, self.pg, self.flat.grad_sym)
        self.autocast = cfg.dtype == "bf16" and cfg.compute == "torch"
        self.step_idx = 0
        self.period = cfg.resolved_sync_period()
        self.model.train()
        self.last_loss = None
        self.last_logits = None
        # seeds differ per rank for dropout etc. after the common initialisation
        torch.manual_seed(cfg.seed + 1000 * (self.rank + 1))
        if cfg.compute == "native":
            from ..ops import native
            native.STEP_STATE["tensor"] = self.exchange.state
        # backward-overlapped bucketed exchange (--overlap): a scheduler fires each bucket's K1+K2 on a side stream
        self.sched = None
        if cfg.overlap and role == "worker" and hasattr(self.exchange, "begin_overlap") and self.period == 1:
            from ..parallel.overlap import BucketScheduler
            self.sched = BucketScheduler(self.flat.params, self.exchange.buckets, self.exchange.bucket_ready).install()
        if role == "worker":
            from ..utils.metrics import logger
            if cfg.overlap and self.sched is None:
                logger.warning("--overlap is implemented for --exchange allgather with sync_period 1; running %s without it",
                               cfg.exchange)
            if cfg.error_feedback and getattr(self.exchange, "residual", None) is None and cfg.resolved_compressor() != "none":
                logger.warning("--error-feedback is implemented for --exchange allgather on the native backend; ignored for %s",
                               cfg.exchange)
        self.graph = None
        self.graph_launches = 0          # launches of this repo's kernels inside one captured step
        self.replays = 0

    # -- CUDA-graph capture of the whole step (forward, backward, exchange, optimizer) ----------
    def capture(self, x: torch.Tensor, y: torch.Tensor, warmup: int = 3) -> None:
        """Capture one full training step in a CUDA graph; later ``step()`` calls copy the batch into the
        static input buffers and replay it.  Every per-step quantity the kernels need (flag sequence
        number, slot parity, RNG position, first-step flag) lives in device memory, so replays are exact.
        Only valid for sync_period == 1 (method 6 alternates two different step bodies)."""
        if self.period != 1:
            raise RuntimeError("graph capture needs sync_period == 1")
        self.static_x, self.static_y = x.clone(), y.clone()
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                self._step_body(self.static_x, self.static_y)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        before = L.launches()
        with torch.cuda.graph(g):
            loss, logits = self._step_body(self.static_x, self.static_y)
        self.graph_launches = L.launches() - before
        self.graph, self.static_loss, self.static_logits = g, loss, logits

    def _forward_backward(self, x, y):
        self.flat.grad.zero_()
        if self.cfg.compute == "native":
            from ..ops import native
            native.begin_step(self.device)
        if self.autocast:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                logits = self.model(x)
            loss = torch.nn.functional.cross_entropy(logits.float(), y)
        else:
            logits = self.model(x)
            loss = Lyr.cross_entropy(logits, y)
        loss.backward()
        if self.cfg.compute == "native":
            native.end_step()
        return loss, logits

    def _overlapped(self, x, y, want_stats: bool = False):
        self.exchange.begin_overlap()
        self.sched.begin_step()
        loss, logits = self._forward_backward(x, y)
        self.sched.flush()
        return loss, logits, self.exchange.finish_overlap(want_stats)

    def _step_body(self, x, y):
        if self.sched is not None:
            loss, logits, _ = self._overlapped(x, y)
            return loss.detach(), logits.detach()
        loss, logits = self._forward_backward(x, y)
        self.exchange.step(self.step_idx, is_sync_step=True, want_stats=False)
        return loss.detach(), logits.detach()

    def own_launches(self) -> int:
        """Launches of this repo's kernels so far (eager launches + replays x launches per captured step)."""
        return L.launches() + self.replays * self.graph_launches

    # -- one optimisation step on a device-resident batch ------------------------------------
    def step(self, x: torch.Tensor, y: torch.Tensor, want_stats: bool = False):
        if self.graph is not None and not want_stats:
            self.static_x.copy_(x, non_blocking=True)
            self.static_y.copy_(y, non_blocking=True)
            self.graph.replay()
            self.replays += 1
            self.step_idx += 1
            self.last_loss, self.last_logits = self.static_loss, self.static_logits
            from ..parallel.host_exchange import ExchangeStats
            return self.static_loss, ExchangeStats()
        if self.sched is not None:
            loss, logits, st = self._overlapped(x, y, want_stats)
        else:
            loss, logits = self._forward_backward(x, y)
            st = self.exchange.step(self.step_idx, is_sync_step=(self.step_idx % self.period == 0), want_stats=want_stats)
        self.step_idx += 1
        self.last_loss, self.last_logits = loss.detach(), logits.detach()
        return self.last_loss, st

    def server_step(self, want_stats: bool = False):
        st = self.exchange.step(self.step_idx, is_sync_step=(self.step_idx % self.period == 0), want_stats=want_stats)
        self.step_idx += 1
        return st

    # -- checkpoint / resume: reference-format state_dict + sidecar with everything a bit-exact resume needs --
    def save_checkpoint(self, path: str, is_writer: bool = True) -> None:
        """``path``: plain ``torch.save(state_dict)`` the reference evaluator loads; ``path + ".state"``: step counter,
        flat momentum buffer, device step state (sequence number / RNG position), error-feedback residual."""
        from ..utils import checkpoint as ckpt
        torch.cuda.synchronize(self.device)
        crc = ckpt.save_model(self.model, path, is_writer)
        extra = {"momentum": self.flat.momentum.detach().cpu() if self.flat.momentum is not None else None,
                 "device_state": self.exchange.state.detach().cpu(),
                 "residual": self.exchange.residual.detach().cpu() if getattr(self.exchange, "residual", None) is not None else None,
                 "adam": {k: v.detach().cpu() for k, v in self.exchange.adam.state_dict().items()}
                 if getattr(self.exchange, "adam", None) is not None else None}
        ckpt.save_sidecar(path, step=self.step_idx, optimizer_state=None, extra=extra, is_writer=is_writer, model_crc=crc,
                          rng_rank=self.rank)

    def load_checkpoint(self, path: str) -> int:
        """Restore weights (+ sidecar when present); returns the step to continue from.  A captured graph stays valid:
        every buffer is updated in place."""
        from ..utils import checkpoint as ckpt
        ckpt.load_model(self.model, path)                 # parameters are views of the flat buffer: copied in place
        if self.flat.param_bf16 is not None:
            self.flat.param_bf16.copy_(self.flat.param)
        side = ckpt.load_sidecar(path)
        if side is not None:
            self.step_idx = int(side["step"])
            ex = side.get("extra") or {}
            if ex.get("momentum") is not None and self.flat.momentum is not None:
                self.flat.momentum.copy_(ex["momentum"])
            if ex.get("device_state") is not None:
                st = ex["device_state"].to(self.exchange.state.device)
                if self.world > 1:
                    # the flag sequence number must stay in step with the peers' mailboxes of THIS job: keep seq, restore the rest
                    st[0] = self.exchange.state[0]
                self.exchange.state.copy_(st)
            if ex.get("residual") is not None and getattr(self.exchange, "residual", None) is not None:
                self.exchange.residual.copy_(ex["residual"])
            if ex.get("adam") is not None and getattr(self.exchange, "adam", None) is not None:
                self.exchange.adam.load_state_dict(ex["adam"])
        torch.cuda.synchronize(self.device)
        return self.step_idx

    def input_spec(self):
        c, h, w = input_shape(self.cfg.network)
        return (self.cfg.batch_size, c, h, w)

    def check_device_error(self) -> int:
        import ctypes as C
        err = C.c_uint32(0)
        L.check(L.load().ewdml_get_device_error(C.byref(err), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        return err.value


def run_training(cfg: TrainConfig, rank: int, world: int, device: torch.device, pg: Optional[PeerGroup] = None,
                 total_steps: Optional[int] = None):
    """The command-line trainer's fast path (``distributed_nn.py --enable-gpu --backend native --graphs``): the loop of
    ``DistributedWorker.train_updated`` / ``SyncReplicasMaster_NN.start_updated`` (``distributed_worker.py:162-239``,
    ``sync_replicas_master_nn.py:158-177``) on :class:`GpuTrainer` -- device-resident batches from :class:`DeviceFeeder`,
    the whole step replayed from one CUDA graph, the loss / accuracy read back only on the steps that are logged,
    reference-format checkpoints every ``eval_freq`` steps.  Returns the list of logged records (workers) or the number of
    served steps (server)."""
    from ..data import DeviceFeeder, build_datasets, dataset_tensors, gpu_augment
    from ..utils.metrics import StepLogger, accuracy, logger
    from ..utils import checkpoint as ckpt
    pg = pg or PeerGroup.from_env(device)
    ps = cfg.exchange == "ps"
    role = "server" if (ps and rank == 0) else "worker"
    tr = GpuTrainer(cfg, device, pg, role=role)
    if cfg.resume:
        tr.load_checkpoint(cfg.resume)
    tr.exchange.broadcast_initial_weights(0)
    n_workers = world - 1 if ps else world
    total = total_steps if total_steps is not None else cfg.max_steps
    will_capture = bool(cfg.graphs) and tr.period == 1 and total - tr.step_idx > 3
    if role == "server":
        # the workers' capture() runs 3 extra (real) warm-up exchanges the step counter does not see
        for _ in range(tr.step_idx, total + (3 if will_capture else 0)):
            tr.server_step()
        torch.cuda.synchronize(device)
        return tr.step_idx
    train_set, _ = build_datasets(cfg)
    data, targets = dataset_tensors(train_set, cfg.dataset)
    shard = (rank - 1) if ps else rank
    feeder = DeviceFeeder(data, targets, cfg.batch_size, device, channels_last=True, seed=cfg.seed,
                          shard_rank=shard if cfg.shard_data else 0, num_shards=n_workers if cfg.shard_data else 1)
    augment = gpu_augment if (cfg.dataset in ("Cifar10", "Cifar100") and not hasattr(train_set, "num_classes")) else None
    log = StepLogger(rank, cfg.metrics_file, cfg.log_interval)
    writer = rank == (1 if ps else 0)
    history = []
    capture_at = tr.step_idx + 3 if will_capture else -1                         # a few eager steps first, then replay
    for step in range(tr.step_idx, total):
        x, y = feeder.next()
        if augment is not None:
            x = augment(x)
        if step == capture_at:
            tr.capture(x, y)                    # runs its own 3 warm-up steps: peers capture at the same step
        loss, st = tr.step(x, y, want_stats=(tr.graph is None and step % max(1, cfg.log_interval) == 0))
        feeder.mark_consumed()
        if step % max(1, cfg.log_interval) == 0:
            p1, p5 = accuracy(tr.last_logits.float(), y, topk=(1, 5))
            history.append(log.log(step, loss=float(loss), prec1=float(p1), prec5=float(p5), bytes_sent=st.bytes_sent,
                                   bytes_recv=st.bytes_recv, dense_bytes=tr.flat.dense_nbytes(), t_send=st.t_send, t_recv=st.t_recv,
                                   images=int(y.shape[0])))
        if step % cfg.eval_freq == 0:
            tr.save_checkpoint(ckpt.model_path(cfg.train_dir, step, cfg.legacy_checkpoint_name), writer)
    torch.cuda.synchronize(device)
    err = tr.check_device_error()
    if err:
        raise RuntimeError(f"device error word {err:#x} (a peer timed out?)")
    log.close()
    return history
