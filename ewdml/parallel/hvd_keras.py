"""Keras-style training loop with Horovod-style callbacks (the capability of ``tensorflow_mnist.py:42-79``).

``tensorflow_mnist.py`` drives ``model.fit`` with ``hvd.DistributedOptimizer(Adam)`` and four callbacks:
``BroadcastGlobalVariablesCallback(0)`` (``:57``), ``MetricAverageCallback()`` (``:63``),
``LearningRateWarmupCallback(initial_lr, warmup_epochs=3)`` (``:68``) and a rank-0 ``ModelCheckpoint(
'./checkpoint-{epoch}.h5')`` (``:72-73``).  There is no TensorFlow here; this module provides the same
pieces on top of :mod:`ewdml.parallel.hvd` for torch modules: a :func:`fit` loop with the Keras callback
protocol (``on_train_begin / on_epoch_begin / on_batch_begin / on_epoch_end``) and the four callbacks.
"""
from __future__ import annotations

import os
from typing import Callable, Dict, Iterable, List, Optional

import torch

from . import hvd


class Callback:
    def set_context(self, model, optimizer):
        self.model, self.optimizer = model, optimizer

    def on_train_begin(self, logs=None): ...
    def on_epoch_begin(self, epoch: int, logs=None): ...
    def on_batch_begin(self, batch: int, logs=None): ...
    def on_batch_end(self, batch: int, logs=None): ...
    def on_epoch_end(self, epoch: int, logs=None): ...


class BroadcastGlobalVariablesCallback(Callback):
    """Every rank starts from rank ``root``'s weights and optimizer state."""

    def __init__(self, root_rank: int = 0):
        self.root = root_rank

    def on_train_begin(self, logs=None):
        hvd.broadcast_parameters(self.model.state_dict(), root_rank=self.root)
        opt = getattr(self.optimizer, "_opt", self.optimizer)
        if opt.state_dict().get("state"):
            hvd.broadcast_optimizer_state(opt, root_rank=self.root)


class MetricAverageCallback(Callback):
    """Averages every numeric entry of ``logs`` over the ranks at the end of an epoch."""

    def on_epoch_end(self, epoch: int, logs=None):
        if logs:
            for k in sorted(logs):
                if isinstance(logs[k], (int, float)):
                    logs[k] = float(hvd.allreduce(torch.tensor(float(logs[k])), name=k))


class LearningRateWarmupCallback(Callback):
    """Horovod's gradual warm-up (Goyal et al. 2017): the learning rate grows from ``initial_lr / size`` to
    ``initial_lr`` over ``warmup_epochs``, adjusted every batch."""

    def __init__(self, initial_lr: float, warmup_epochs: float = 5, steps_per_epoch: Optional[int] = None, verbose: int = 0):
        self.initial_lr, self.warmup_epochs, self.steps_per_epoch, self.verbose = initial_lr, warmup_epochs, steps_per_epoch, verbose
        self._epoch = 0

    def multiplier(self, epoch: float) -> float:
        size = hvd.size()
        if epoch >= self.warmup_epochs or self.warmup_epochs <= 0:
            return 1.0
        return 1.0 / size * (epoch * (size - 1) / self.warmup_epochs + 1)

    def on_epoch_begin(self, epoch: int, logs=None):
        self._epoch = epoch

    def on_batch_begin(self, batch: int, logs=None):
        spe = self.steps_per_epoch or (logs or {}).get("steps_per_epoch", 1)
        lr = self.initial_lr * self.multiplier(self._epoch + batch / max(1, spe))
        for g in self.optimizer.param_groups:
            g["lr"] = lr

    def on_epoch_end(self, epoch: int, logs=None):
        if self.verbose and hvd.rank() == 0 and epoch + 1 == int(self.warmup_epochs):
            print("Epoch %d: finished gradual learning rate warmup to %g." % (epoch + 1, self.initial_lr))


class ModelCheckpoint(Callback):
    """``state_dict`` per epoch (``{epoch}`` is 1-based as in Keras); add it on rank 0 only, like the reference does."""

    def __init__(self, filepath: str):
        self.filepath = filepath

    def on_epoch_end(self, epoch: int, logs=None):
        path = self.filepath.format(epoch=epoch + 1, **(logs or {}))
        tmp = path + ".tmp"
        torch.save(self.model.state_dict(), tmp)
        os.replace(tmp, path)


def fit(model: torch.nn.Module, optimizer, loss_fn: Callable, batches: Iterable, steps_per_epoch: int, epochs: int,
        callbacks: Optional[List[Callback]] = None, device: Optional[torch.device] = None, verbose: int = 0) -> List[Dict]:
    """``model.fit(dataset, steps_per_epoch=..., epochs=..., callbacks=...)``; ``batches`` is an endless iterator of
    ``(x, y)`` (``dataset.repeat()``).  Returns the per-epoch logs (``loss``, ``accuracy``)."""
    callbacks = callbacks or []
    for cb in callbacks:
        cb.set_context(model, optimizer)
        cb.on_train_begin()
    history = []
    it = iter(batches)
    for epoch in range(epochs):
        model.train()
        for cb in callbacks:
            cb.on_epoch_begin(epoch)
        loss_sum, correct, seen = 0.0, 0, 0
        for b in range(steps_per_epoch):
            for cb in callbacks:
                cb.on_batch_begin(b, {"steps_per_epoch": steps_per_epoch})
            x, y = next(it)
            if device is not None:
                x, y = x.to(device), y.to(device)
            optimizer.zero_grad()
            out = model(x)
            loss = loss_fn(out, y)
            loss.backward()
            if hasattr(optimizer, "synchronize"):
                optimizer.synchronize()
            optimizer.step()
            loss_sum += float(loss.detach()) * y.shape[0]
            correct += int((out.argmax(1) == y).sum())
            seen += int(y.shape[0])
            for cb in callbacks:
                cb.on_batch_end(b)
        logs = {"loss": loss_sum / max(1, seen), "accuracy": correct / max(1, seen)}
        for cb in callbacks:
            cb.on_epoch_end(epoch, logs)
        if verbose:
            print("Epoch %d/%d - loss: %.4f - accuracy: %.4f" % (epoch + 1, epochs, logs["loss"], logs["accuracy"]))
        history.append(logs)
    return history
