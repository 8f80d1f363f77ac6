"""Polling evaluator protocol (src/distributed_evaluator.py:72-110)."""
import threading
import time

import torch

from ewdml.config import TrainConfig
from ewdml.data import prepare_data
from ewdml.models import build_model
from ewdml.train.evaluator import DistributedEvaluator
from ewdml.utils import checkpoint as ckpt


def test_evaluator_polls_and_evaluates(tmp_path):
    cfg = TrainConfig(dataset="MNIST", test_batch_size=64, data="synthetic", synthetic_size=256)
    _, test_loader = prepare_data(cfg)
    d = str(tmp_path) + "/"
    ev = DistributedEvaluator("LeNet", d, eval_freq=5, eval_batch_size=64, poll_s=0.05)

    def writer():
        time.sleep(0.2)
        ckpt.save_model(build_model("LeNet"), ckpt.model_path(d))
        time.sleep(0.3)
        ckpt.save_model(build_model("LeNet"), ckpt.model_path(d))

    t = threading.Thread(target=writer); t.start()
    res = ev.evaluate(test_loader, max_evals=2, timeout=20)
    t.join()
    assert [r["step"] for r in res] == [5, 10]
    assert all(0 <= r["accuracy"] <= 100 and r["total"] == 64 for r in res)


def test_evaluator_times_out_without_checkpoint(tmp_path):
    cfg = TrainConfig(dataset="MNIST", test_batch_size=8, data="synthetic", synthetic_size=32)
    _, test_loader = prepare_data(cfg)
    ev = DistributedEvaluator("LeNet", str(tmp_path) + "/", poll_s=0.05)
    assert ev.evaluate(test_loader, max_evals=1, timeout=0.3) == []
