"""Per-rank launcher for the UNMODIFIED reference entry point (baseline/_ref/distributed_nn.py).

The reference downloads MNIST/CIFAR through torchvision; the image blobs are not in the reference
checkout and there is no network, so -- and only so -- ``torchvision.datasets.{MNIST,CIFAR10,
CIFAR100,SVHN}`` are replaced by synthetic datasets of the same shape before the reference module
is executed with ``runpy`` (its code, flags, model, compressor objects, optimizer and gloo
collectives are untouched).  A logging handler timestamps the reference's own per-step log line
("Worker: r, Step: s, ...", distributed_worker.py:230-231) so throughput can be computed without
editing it.
"""
import json
import logging
import os
import runpy
import sys
import time

import torch
import torchvision

REF = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


class _Fake(torch.utils.data.Dataset):
    shape, classes = (3, 32, 32), 10

    def __init__(self, root=None, train=True, split=None, download=False, transform=None, **kw):
        n = int(os.environ.get("EWDML_REF_DATASET_SIZE", "4096"))
        g = torch.Generator().manual_seed(1234 + (0 if train and split != "test" else 1))
        self.data = torch.randn(n, *self.shape, generator=g)
        self.targets = torch.randint(0, self.classes, (n,), generator=g)

    def __len__(self):
        return self.data.shape[0]

    def __getitem__(self, i):
        return self.data[i], int(self.targets[i])


class FakeMNIST(_Fake):
    shape = (1, 28, 28)


class FakeCIFAR10(_Fake):
    pass


class FakeCIFAR100(_Fake):
    classes = 100


class StepTimer(logging.Handler):
    def __init__(self):
        super().__init__()
        self.stamps = {}

    def emit(self, record):
        msg = record.getMessage()
        if msg.startswith("Worker:") and "Step:" in msg:
            try:
                step = int(msg.split("Step:")[1].split(",")[0])
                self.stamps.setdefault(step, time.perf_counter())
            except Exception:
                pass


def main():
    torchvision.datasets.MNIST = FakeMNIST
    torchvision.datasets.CIFAR10 = FakeCIFAR10
    torchvision.datasets.CIFAR100 = FakeCIFAR100
    torchvision.datasets.SVHN = FakeCIFAR10
    timer = StepTimer()
    logging.getLogger().addHandler(timer)
    sys.path.insert(0, REF)
    out = os.environ.get("EWDML_REF_OUT")
    rank = int(os.environ["RANK"])
    t0 = time.perf_counter()
    status = "ok"
    try:
        runpy.run_path(os.path.join(REF, "distributed_nn.py"), run_name="__main__")
    except SystemExit as e:
        status = f"exit {e.code}"
    except BaseException as e:  # noqa: BLE001 -- report, the parent decides
        status = f"{type(e).__name__}: {e}"
    if out:
        with open(os.path.join(out, f"rank{rank}.json"), "w") as f:
            json.dump({"rank": rank, "status": status, "t0": t0, "stamps": timer.stamps}, f)
    if status != "ok":
        sys.stderr.write(f"[ref_shim] rank {rank}: {status}\n")
        os._exit(1)


if __name__ == "__main__":
    main()
