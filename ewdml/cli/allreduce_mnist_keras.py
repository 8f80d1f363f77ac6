"""Allreduce MNIST trainer in the shape of the reference's Keras script (``tensorflow_mnist.py:1-79``).

    python -m ewdml.cli.launch -np 4 -- python -m ewdml.cli.allreduce_mnist_keras --epochs 2

Same network (Conv 32 / Conv 64 / MaxPool / Dropout .25 / Dense 128 / Dropout .5 / Dense 10, ``:26-35``), Adam with
``lr = 0.001 * size`` (``:38-39``) wrapped in ``DistributedOptimizer`` (``:42-43``), the four callbacks
(``:53-73``) and ``steps_per_epoch = 500 // size`` (``:79``).  No compression, as in the reference.
"""
from __future__ import annotations

import argparse
import itertools

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..config import TrainConfig
from ..data import build_datasets
from ..parallel import hvd
from ..parallel import hvd_keras as K


class KerasMnistNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 32, 3)
        self.conv2 = nn.Conv2d(32, 64, 3)
        self.drop1, self.drop2 = nn.Dropout(0.25), nn.Dropout(0.5)
        self.fc1 = nn.Linear(64 * 12 * 12, 128)
        self.fc2 = nn.Linear(128, 10)

    def forward(self, x):
        x = F.relu(self.conv2(F.relu(self.conv1(x))))
        x = self.drop1(F.max_pool2d(x, 2))
        x = self.drop2(F.relu(self.fc1(torch.flatten(x, 1))))
        return self.fc2(x)               # logits; the loss applies the softmax (SparseCategoricalCrossentropy)


def main(argv=None):
    p = argparse.ArgumentParser(description="ewdml allreduce MNIST example, Keras-style loop")
    p.add_argument("--batch-size", type=int, default=128)
    p.add_argument("--epochs", type=int, default=24)
    p.add_argument("--lr", type=float, default=0.001)
    p.add_argument("--steps-per-epoch", type=int, default=500, help="divided by the number of ranks")
    p.add_argument("--warmup-epochs", type=float, default=3)
    p.add_argument("--checkpoint-format", default="./checkpoint-{epoch}.pt")
    p.add_argument("--no-cuda", action="store_true", default=False)
    p.add_argument("--seed", type=int, default=42)
    p.add_argument("--synthetic-size", type=int, default=4096)
    p.add_argument("--data-dir", default=None)
    a = p.parse_args(argv)

    cuda = not a.no_cuda and torch.cuda.is_available()
    hvd.init("nccl" if cuda else "gloo")
    torch.manual_seed(a.seed + hvd.rank())
    device = torch.device("cuda", hvd.local_rank() % torch.cuda.device_count()) if cuda else torch.device("cpu")
    torch.set_num_threads(1)
    cfg = TrainConfig(dataset="MNIST", batch_size=a.batch_size, data_root=a.data_dir or ".", synthetic_size=a.synthetic_size)
    train_set, _ = build_datasets(cfg)
    sampler = torch.utils.data.distributed.DistributedSampler(train_set, num_replicas=hvd.size(), rank=hvd.rank())
    loader = torch.utils.data.DataLoader(train_set, batch_size=a.batch_size, sampler=sampler, drop_last=True)

    def repeat():                         # dataset.repeat().shuffle(...)
        for epoch in itertools.count():
            sampler.set_epoch(epoch)
            yield from loader

    model = KerasMnistNet().to(device)
    scaled_lr = a.lr * hvd.size()
    opt = hvd.DistributedOptimizer(torch.optim.Adam(model.parameters(), lr=scaled_lr), named_parameters=model.named_parameters())
    callbacks = [K.BroadcastGlobalVariablesCallback(0), K.MetricAverageCallback(),
                 K.LearningRateWarmupCallback(initial_lr=scaled_lr, warmup_epochs=a.warmup_epochs, verbose=1)]
    if hvd.rank() == 0 and a.checkpoint_format:
        callbacks.append(K.ModelCheckpoint(a.checkpoint_format))
    hist = K.fit(model, opt, F.cross_entropy, repeat(), steps_per_epoch=max(1, a.steps_per_epoch // hvd.size()),
                 epochs=a.epochs, callbacks=callbacks, device=device, verbose=1 if hvd.rank() == 0 else 0)
    hvd.shutdown()
    return hist


if __name__ == "__main__":
    main()
