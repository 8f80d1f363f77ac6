"""``distributed_evaluator.py`` entry point (flags of ``src/distributed_evaluator.py:37-54``)."""
from __future__ import annotations

import argparse

import torch

from ..config import TrainConfig
from ..data import prepare_data
from ..train.evaluator import DistributedEvaluator


def main(argv=None):
    p = argparse.ArgumentParser(description="ewdml polling evaluator")
    p.add_argument("--eval-batch-size", type=int, default=10000)
    p.add_argument("--eval-freq", type=int, default=50)
    p.add_argument("--model-dir", type=str, default="output/models/")
    p.add_argument("--dataset", type=str, default="MNIST")
    p.add_argument("--network", type=str, default="LeNet")
    p.add_argument("--data", type=str, default="auto")
    p.add_argument("--data-root", type=str, default=".")
    p.add_argument("--poll-seconds", type=float, default=10.0)
    p.add_argument("--max-evals", type=int, default=0, help="0 = loop forever like the reference")
    p.add_argument("--timeout", type=float, default=0.0)
    p.add_argument("--enable-gpu", action="store_true")
    a = p.parse_args(argv)
    cfg = TrainConfig(dataset=a.dataset, network=a.network, test_batch_size=a.eval_batch_size,
                      data=a.data, data_root=a.data_root).validate()
    _, test_loader = prepare_data(cfg)
    dev = torch.device("cuda") if (a.enable_gpu and torch.cuda.is_available()) else torch.device("cpu")
    ev = DistributedEvaluator(a.network, a.model_dir, a.eval_freq, a.eval_batch_size,
                              num_classes=cfg.num_classes(), device=dev, poll_s=a.poll_seconds)
    ev.evaluate(test_loader, max_evals=a.max_evals or None, timeout=a.timeout or None)


if __name__ == "__main__":
    main()
