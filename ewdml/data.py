"""Datasets and loaders.

Capability parity with ``prepare_data`` (``PyTorch-parameter-server/src/util.py:20-106``): MNIST,
Cifar10, Cifar100 and SVHN with the same normalisation constants and augmentation, when the real
image blobs are on disk (``--data real``).  The build box has no datasets and no network, so the
default (``--data auto``) falls back to a *learnable synthetic* set of the same shape: class
conditional patterns plus noise from a fixed teacher, so "accuracy after N steps" is a meaningful
number.

Fixes relative to the reference: every worker gets its own shard
(``shard_data``; the reference iterates the full set on every worker, see the commented stub at
``distributed_worker.py:175-181``) and shuffling is seeded.

The GPU path does not use a Python ``DataLoader`` at all: :class:`DeviceFeeder` keeps the (small)
dataset in pinned host memory and issues one ``cudaMemcpyAsync`` per batch on a copy stream, one
step ahead of compute.
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass
from typing import Iterator, Optional, Tuple

import torch

# (channels, height, width, classes, mean, std)
_SHAPES = {
    "MNIST": (1, 28, 28, 10, (0.1307,), (0.3081,)),
    "Cifar10": (3, 32, 32, 10, tuple(x / 255.0 for x in (125.3, 123.0, 113.9)),
                tuple(x / 255.0 for x in (63.0, 62.1, 66.7))),
    "Cifar100": (3, 32, 32, 100, tuple(x / 255.0 for x in (125.3, 123.0, 113.9)),
                 tuple(x / 255.0 for x in (63.0, 62.1, 66.7))),
    "SVHN": (3, 32, 32, 10, (0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010)),
}


def dataset_shape(name: str) -> Tuple[int, int, int, int]:
    c, h, w, k = _SHAPES[name][:4]
    return c, h, w, k


class SyntheticImageDataset(torch.utils.data.Dataset):
    """Class-conditional patterns + Gaussian noise, already normalised.

    ``x = pattern[y] * signal + noise`` with one fixed low-frequency random pattern per class.
    A linear probe reaches ~100 % on it, a randomly initialised CNN starts at chance, so
    convergence differences between compressors show up within tens of steps.
    """

    def __init__(self, name: str, n: int, train: bool = True, seed: int = 1234,
                 signal: float = 0.6) -> None:
        c, h, w, k = dataset_shape(name)
        g = torch.Generator().manual_seed(seed)
        # low-frequency class templates (upsampled 7x7 / 8x8 random fields)
        base = torch.randn(k, c, max(1, h // 4), max(1, w // 4), generator=g)
        templ = torch.nn.functional.interpolate(base, size=(h, w), mode="bilinear",
                                                align_corners=False)
        g2 = torch.Generator().manual_seed(seed + (1 if train else 2))
        self.targets = torch.randint(0, k, (n,), generator=g2)
        noise = torch.randn(n, c, h, w, generator=g2)
        self.data = (templ[self.targets] * signal + noise).contiguous()
        self.num_classes = k

    def __len__(self) -> int:
        return self.data.shape[0]

    def __getitem__(self, i: int):
        return self.data[i], self.targets[i]


def _real_datasets(cfg):
    """torchvision datasets with the reference's transforms (util.py:22-104)."""
    from torchvision import datasets, transforms
    import torch.nn.functional as F

    name = cfg.dataset
    _, _, _, _, mean, std = _SHAPES[name]
    norm = transforms.Normalize(mean, std)
    root = cfg.data_root
    if name == "MNIST":
        tf = transforms.Compose([transforms.ToTensor(), norm])
        return (datasets.MNIST(os.path.join(root, "mnist_data"), train=True, download=False, transform=tf),
                datasets.MNIST(os.path.join(root, "mnist_data"), train=False, download=False, transform=tf))
    if name in ("Cifar10", "Cifar100"):
        reflect = transforms.Lambda(
            lambda x: F.pad(x.unsqueeze(0), (4, 4, 4, 4), mode="reflect").squeeze(0))
        tf_train = transforms.Compose([transforms.ToTensor(), reflect, transforms.ToPILImage(),
                                       transforms.RandomCrop(32), transforms.RandomHorizontalFlip(),
                                       transforms.ToTensor(), norm])
        tf_test = transforms.Compose([transforms.ToTensor(), norm])
        cls = datasets.CIFAR10 if name == "Cifar10" else datasets.CIFAR100
        sub = "cifar10_data" if name == "Cifar10" else "cifar100_data"
        return (cls(os.path.join(root, sub), train=True, download=False, transform=tf_train),
                cls(os.path.join(root, sub), train=False, download=False, transform=tf_test))
    if name == "SVHN":
        tf_train = transforms.Compose([transforms.RandomCrop(32, padding=4),
                                       transforms.RandomHorizontalFlip(), transforms.ToTensor(), norm])
        tf_test = transforms.Compose([transforms.ToTensor(), norm])
        return (datasets.SVHN(os.path.join(root, "svhn_data"), split="train", download=False, transform=tf_train),
                datasets.SVHN(os.path.join(root, "svhn_data"), split="test", download=False, transform=tf_test))
    raise ValueError(name)


def build_datasets(cfg):
    """Return (train_set, test_set); synthetic unless real blobs are present/forced."""
    if cfg.data in ("real", "auto"):
        try:
            return _real_datasets(cfg)
        except Exception:
            if cfg.data == "real":
                raise
    n_train = cfg.synthetic_size
    return (SyntheticImageDataset(cfg.dataset, n_train, train=True, seed=1234),
            SyntheticImageDataset(cfg.dataset, max(cfg.test_batch_size, n_train // 4), train=False,
                                  seed=1234))


def dataset_tensors(ds, name: str) -> Tuple[torch.Tensor, torch.Tensor]:
    """(normalised fp32 NCHW images, int64 labels) of a whole dataset, for :class:`DeviceFeeder`.  Synthetic sets are stored
    that way already; torchvision sets are converted from their raw uint8 arrays with the reference's constants
    (``util.py:22-104``) -- augmentation then happens on the device (:func:`gpu_augment`)."""
    if isinstance(ds, SyntheticImageDataset):
        return ds.data, ds.targets
    import numpy as np
    raw = np.asarray(ds.data)
    x = torch.from_numpy(raw)
    if x.dim() == 3:                       # MNIST [N, H, W]
        x = x.unsqueeze(1)
    elif x.shape[-1] in (1, 3):            # CIFAR [N, H, W, C]; SVHN is already [N, C, H, W]
        x = x.permute(0, 3, 1, 2)
    _, _, _, _, mean, std = _SHAPES[name]
    x = x.float().div_(255.0)
    x = (x - torch.tensor(mean).view(1, -1, 1, 1)) / torch.tensor(std).view(1, -1, 1, 1)
    labels = getattr(ds, "targets", None)
    if labels is None:
        labels = ds.labels
    return x.contiguous(), torch.as_tensor(np.asarray(labels), dtype=torch.int64)


def gpu_augment(x: torch.Tensor, pad: int = 4) -> torch.Tensor:
    """The reference's CIFAR training augmentation (reflect-pad 4, random 32x32 crop, random horizontal flip,
    ``util.py:36-47``) applied to a whole device batch: one crop offset and one flip decision per image."""
    n, c, h, w = x.shape
    xp = torch.nn.functional.pad(x, (pad, pad, pad, pad), mode="reflect")
    dy = torch.randint(0, 2 * pad + 1, (n,), device=x.device)
    dx = torch.randint(0, 2 * pad + 1, (n,), device=x.device)
    flip = torch.rand(n, device=x.device) < 0.5
    rows = (dy.view(n, 1) + torch.arange(h, device=x.device).view(1, h))                     # [n, h]
    cols = (dx.view(n, 1) + torch.arange(w, device=x.device).view(1, w))                     # [n, w]
    cols = torch.where(flip.view(n, 1), cols.flip(1), cols)
    idx = (rows.view(n, h, 1) * (w + 2 * pad) + cols.view(n, 1, w)).view(n, 1, h * w).expand(n, c, h * w)
    return xp.reshape(n, c, -1).gather(2, idx).view(n, c, h, w).contiguous(memory_format=torch.channels_last)


def prepare_data(cfg, shard_rank: int = 0, num_shards: int = 1):
    """(train_loader, test_loader) — the ``prepare_data(args)`` entry point of util.py:20.

    ``shard_rank``/``num_shards`` select this worker's partition (DistributedSampler semantics,
    as in ``horvod_pytorch.py:154-157``); pass 0/1 for the reference's unsharded behaviour.
    """
    train_set, test_set = build_datasets(cfg)
    gen = torch.Generator().manual_seed(cfg.seed)
    sampler = None
    if num_shards > 1 and cfg.shard_data:
        sampler = torch.utils.data.distributed.DistributedSampler(
            train_set, num_replicas=num_shards, rank=shard_rank, shuffle=True, seed=cfg.seed)
    train_loader = torch.utils.data.DataLoader(
        train_set, batch_size=cfg.batch_size, shuffle=(sampler is None), sampler=sampler,
        generator=gen, drop_last=True)
    test_loader = torch.utils.data.DataLoader(test_set, batch_size=cfg.test_batch_size, shuffle=False)
    return train_loader, test_loader


@dataclass
class _Slot:
    x: torch.Tensor
    y: torch.Tensor
    ev: "torch.cuda.Event"


class DeviceFeeder:
    """Pinned-host -> device batch pipeline for the GPU path (double-buffered, copy stream).

    Holds the (sharded) dataset in host memory in NHWC order.  ``next()`` returns the batch whose H2D copy was
    enqueued earlier; ``mark_consumed()`` -- called right after the step has been launched -- gathers the
    FOLLOWING shuffled batch into a pinned staging buffer and enqueues its copy on a side stream, so both
    the host-side gather and the copy overlap the GPU work of the current step.
    Batches are returned as NCHW-shaped channels-last views.  ``h2d_bytes_per_step`` is what
    bench.py reports.
    """

    def __init__(self, data: torch.Tensor, targets: torch.Tensor, batch_size: int,
                 device: torch.device, dtype: torch.dtype = torch.float32,
                 channels_last: bool = True, seed: int = 1, shard_rank: int = 0,
                 num_shards: int = 1) -> None:
        n = data.shape[0]
        idx = torch.arange(shard_rank, n, num_shards)
        self.nhwc = bool(channels_last and data.dim() == 4)
        d = data[idx].to(dtype)
        self.data = (d.permute(0, 2, 3, 1) if self.nhwc else d).contiguous()
        self.targets = targets[idx].to(torch.int64).contiguous()
        self.n = self.data.shape[0]
        self.bs = batch_size
        self.device = device
        self.gen = torch.Generator().manual_seed(seed + 7919 * shard_rank)
        self.steps_per_epoch = max(1, self.n // batch_size)
        self._pos = 0
        self._perm = torch.randperm(self.n, generator=self.gen)
        self.h2d_bytes_per_step = batch_size * (self.data[0].numel() * self.data.element_size()
                                                + self.targets.element_size())
        self._cuda = device.type == "cuda"
        if self._cuda:
            self._stream = torch.cuda.Stream(device=device)
            shape = (batch_size,) + tuple(self.data.shape[1:])
            self._hx = [torch.empty(shape, dtype=dtype).pin_memory() for _ in range(2)]
            self._hy = [torch.empty(batch_size, dtype=torch.int64).pin_memory() for _ in range(2)]
            self._dx = [torch.empty(shape, dtype=dtype, device=device) for _ in range(2)]
            self._dy = [torch.empty(batch_size, dtype=torch.int64, device=device) for _ in range(2)]
            self._ev = [torch.cuda.Event() for _ in range(2)]
            self._consumed = [None, None]
            self._ready = [False, False]
            self._k, self._last = 0, 0
            self._prefetch(0)

    def _next_indices(self) -> torch.Tensor:
        if self._pos + self.bs > self.n:
            self._perm = torch.randperm(self.n, generator=self.gen)
            self._pos = 0
        sel = self._perm[self._pos:self._pos + self.bs]
        self._pos += self.bs
        return sel

    def _as_nchw(self, x: torch.Tensor) -> torch.Tensor:
        return x.permute(0, 3, 1, 2) if self.nhwc else x

    def _prefetch(self, k: int) -> None:
        """Gather the next shuffled batch into pinned slot k and enqueue its H2D copy on the copy stream."""
        sel = self._next_indices()
        self._ev[k].synchronize()                      # the previous copy out of the pinned staging buffer k has finished
        if self._consumed[k] is not None:
            self._stream.wait_event(self._consumed[k])  # the step that read device buffer k has finished
        torch.index_select(self.data, 0, sel, out=self._hx[k])
        torch.index_select(self.targets, 0, sel, out=self._hy[k])
        with torch.cuda.stream(self._stream):
            self._dx[k].copy_(self._hx[k], non_blocking=True)
            self._dy[k].copy_(self._hy[k], non_blocking=True)
            self._ev[k].record(self._stream)
        self._ready[k] = True

    def next(self) -> Tuple[torch.Tensor, torch.Tensor]:
        """Device batch for this step.  Call ``mark_consumed()`` right after launching the step: the host-side
        gather + H2D copy of the FOLLOWING batch then overlaps the GPU work of this one."""
        if not self._cuda:
            sel = self._next_indices()
            return self._as_nchw(self.data[sel]), self.targets[sel]
        k = self._k
        if not self._ready[k]:
            self._prefetch(k)
        torch.cuda.current_stream(self.device).wait_event(self._ev[k])
        self._ready[k] = False
        self._k ^= 1
        self._last = k
        return self._as_nchw(self._dx[k]), self._dy[k]

    def mark_consumed(self) -> None:
        """The step using the batch from the last ``next()`` has been enqueued: remember when it finishes and start
        preparing the next batch now, while the GPU is busy."""
        if not self._cuda:
            return
        k = self._last
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self._consumed[k] = ev
        if not self._ready[self._k]:
            self._prefetch(self._k)

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        for _ in range(self.steps_per_epoch):
            yield self.next()
