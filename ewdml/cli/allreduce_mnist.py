"""Allreduce MNIST trainer: the capability of ``horvod_pytorch.py`` (and of the Keras variant
``tensorflow_mnist.py``: LR warm-up, metric averaging, rank-0 per-epoch checkpoints) without Horovod.

    python -m ewdml.cli.launch -np 4 -- python -m ewdml.cli.allreduce_mnist --epochs 1

Flags follow ``horvod_pytorch.py:16-40``.  Data is sharded with a DistributedSampler, the learning
rate is scaled by the number of ranks (not for Adasum), parameters and optimizer state are broadcast
from rank 0, gradients go through ``DistributedOptimizer(compression=QSGD)``.
"""
from __future__ import annotations

import argparse
import os

import torch
import torch.nn.functional as F

from ..config import TrainConfig
from ..data import build_datasets
from ..models import MnistNet
from ..parallel import hvd
from ..compress import Compression


def main(argv=None):
    p = argparse.ArgumentParser(description="ewdml allreduce MNIST example")
    p.add_argument("--batch-size", type=int, default=64)
    p.add_argument("--test-batch-size", type=int, default=1000)
    p.add_argument("--epochs", type=int, default=1)
    p.add_argument("--lr", type=float, default=0.01)
    p.add_argument("--momentum", type=float, default=0.5)
    p.add_argument("--no-cuda", action="store_true", default=False)
    p.add_argument("--seed", type=int, default=42)
    p.add_argument("--log-interval", type=int, default=10)
    p.add_argument("--fp16-allreduce", action="store_true", default=False)
    p.add_argument("--use-adasum", action="store_true", default=False)
    p.add_argument("--gradient-predivide-factor", type=float, default=1.0)
    p.add_argument("--data-dir", default=None)
    p.add_argument("--compression", default="qsgd", choices=["none", "fp16", "qsgd"])
    p.add_argument("--warmup-epochs", type=float, default=0.0, help="LR warm-up (tensorflow_mnist.py:61-66)")
    p.add_argument("--checkpoint-format", default="", help="rank-0 checkpoint per epoch, e.g. ./checkpoint-{epoch}.pt")
    p.add_argument("--max-steps", type=int, default=0)
    p.add_argument("--synthetic-size", type=int, default=2048)
    p.add_argument("--backend", default="auto", choices=["auto", "native", "nccl", "gloo"],
                   help="native = this framework's NVLink kernels (K1/K2 for compressed gradients, the collective engine for the rest); "
                        "auto = native on CUDA when libewdml.so is built, else nccl / gloo")
    a = p.parse_args(argv)

    cuda = not a.no_cuda and torch.cuda.is_available()
    backend = a.backend
    if backend == "auto":
        from ..runtime import lib as _L
        backend = ("native" if _L.available() else "nccl") if cuda else "gloo"
    hvd.init(backend)
    torch.manual_seed(a.seed)
    device = torch.device("cuda", hvd.local_rank() % torch.cuda.device_count()) if cuda else torch.device("cpu")
    if cuda:
        torch.cuda.set_device(device)
    torch.set_num_threads(1)
    cfg = TrainConfig(dataset="MNIST", batch_size=a.batch_size, test_batch_size=a.test_batch_size,
                      data_root=a.data_dir or ".", synthetic_size=a.synthetic_size)
    train_set, test_set = build_datasets(cfg)
    train_sampler = torch.utils.data.distributed.DistributedSampler(train_set, num_replicas=hvd.size(), rank=hvd.rank())
    test_sampler = torch.utils.data.distributed.DistributedSampler(test_set, num_replicas=hvd.size(), rank=hvd.rank())
    train_loader = torch.utils.data.DataLoader(train_set, batch_size=a.batch_size, sampler=train_sampler)
    test_loader = torch.utils.data.DataLoader(test_set, batch_size=a.test_batch_size, sampler=test_sampler)

    model = MnistNet().to(device)
    lr_scaler = hvd.size() if not a.use_adasum else 1
    base_lr = a.lr * lr_scaler
    optimizer = torch.optim.SGD(model.parameters(), lr=base_lr, momentum=a.momentum)
    hvd.broadcast_parameters(model.state_dict(), root_rank=0)
    hvd.broadcast_optimizer_state(optimizer, root_rank=0)
    compression = {"none": Compression.none, "fp16": Compression.fp16, "qsgd": Compression.qsgd}[
        "fp16" if a.fp16_allreduce else a.compression]
    optimizer = hvd.DistributedOptimizer(optimizer, named_parameters=model.named_parameters(), compression=compression,
                                         op=hvd.Adasum if a.use_adasum else hvd.Average,
                                         gradient_predivide_factor=a.gradient_predivide_factor)

    def metric_average(val, name):
        return hvd.allreduce(torch.tensor(float(val)), name=name).item()

    steps = 0
    warm_steps = int(a.warmup_epochs * len(train_loader))
    for epoch in range(1, a.epochs + 1):
        model.train()
        train_sampler.set_epoch(epoch)
        for batch_idx, (data, target) in enumerate(train_loader):
            if warm_steps and steps < warm_steps:       # linear warm-up from lr/size to lr
                for g in optimizer.param_groups:
                    g["lr"] = a.lr + (base_lr - a.lr) * steps / warm_steps
            data, target = data.to(device), target.to(device)
            optimizer.zero_grad()
            loss = F.nll_loss(model(data), target)
            loss.backward()
            optimizer.synchronize()
            optimizer.step()
            steps += 1
            if batch_idx % a.log_interval == 0 and hvd.rank() == 0:
                print("Train Epoch: {} [{}/{} ({:.0f}%)]\tLoss: {:.6f}".format(
                    epoch, batch_idx * len(data), len(train_sampler), 100.0 * batch_idx / len(train_loader), loss.item()))
            if a.max_steps and steps >= a.max_steps:
                break
        model.eval()
        test_loss, test_acc = 0.0, 0.0
        with torch.no_grad():
            for data, target in test_loader:
                data, target = data.to(device), target.to(device)
                out = model(data)
                test_loss += F.nll_loss(out, target, reduction="sum").item()
                test_acc += out.argmax(1).eq(target).float().sum().item()
        test_loss = metric_average(test_loss / len(test_sampler), "avg_loss")
        test_acc = metric_average(test_acc / len(test_sampler), "avg_accuracy")
        if hvd.rank() == 0:
            print("\nTest set: Average loss: {:.4f}, Accuracy: {:.2f}%\n".format(test_loss, 100.0 * test_acc))
            if a.checkpoint_format:
                torch.save(model.state_dict(), a.checkpoint_format.format(epoch=epoch))
        if a.max_steps and steps >= a.max_steps:
            break
    hvd.shutdown()


if __name__ == "__main__":
    main()
