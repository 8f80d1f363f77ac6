"""Stand-alone polling evaluator (``src/distributed_evaluator.py:56-141``).

Protocol kept from the reference: the filesystem is the process boundary -- a trainer overwrites
``model_dir + "model_step_"`` every ``eval_freq`` steps and this process polls for it
(``:109-110``), loads the ``state_dict`` into a freshly built network (``:104-107``) and prints
average NLL + accuracy over the test set (``:89-103``); if the file is missing it sleeps
(10 s in the reference, configurable here) and retries.

Added: ``max_evals`` / ``timeout`` so the loop can terminate (the reference loops forever), and
the file's mtime is tracked so the same checkpoint is not evaluated twice while the trainer is
between saves.
"""
from __future__ import annotations

import os
import time
from typing import Optional

import torch
import torch.nn.functional as F

from ..models import build_model
from ..utils.checkpoint import model_path


class DistributedEvaluator:
    def __init__(self, network: str, model_dir: str, eval_freq: int = 50, eval_batch_size: int = 10000,
                 num_classes: int = 10, device: Optional[torch.device] = None, poll_s: float = 10.0):
        self._cur_step = 0
        self._model_dir = model_dir
        self._eval_freq = int(eval_freq)
        self._eval_batch_size = eval_batch_size
        self.network_config = network
        self.num_classes = num_classes
        self.device = device or torch.device("cpu")
        self.poll_s = poll_s
        self.results = []
        self._last_mtime = None

    def _model_dir_generator(self, next_step_to_fetch: int) -> str:
        return model_path(self._model_dir, None, True)

    def _load_model(self, file_path: str):
        self.network = build_model(self.network_config, num_classes=self.num_classes)
        with open(file_path, "rb") as f:
            self.network.load_state_dict(torch.load(f, map_location="cpu", weights_only=True))
        self.network.to(self.device)

    @torch.no_grad()
    def _evaluate_model(self, test_loader):
        self.network.eval()
        loss, correct, n = 0.0, 0, 0
        for data, target in test_loader:
            data, target = data.to(self.device), target.to(self.device)
            out = self.network(data).float()
            lp = out if getattr(self.network, "log_softmax_output", False) else F.log_softmax(out, dim=1)
            loss += F.nll_loss(lp, target, reduction="sum").item()
            correct += int(out.argmax(1).eq(target).sum())
            n += target.shape[0]
        loss /= max(1, n)
        print("\nTest set: Avg. loss: {:.4f}, Accuracy: {}/{} ({:.0f}%)\n".format(
            loss, correct, n, 100.0 * correct / max(1, n)))
        return {"loss": loss, "correct": correct, "total": n, "accuracy": 100.0 * correct / max(1, n)}

    def evaluate(self, validation_loader, max_evals: Optional[int] = None, timeout: Optional[float] = None):
        self._next_step_to_fetch = self._cur_step + self._eval_freq
        deadline = None if timeout is None else time.time() + timeout
        while max_evals is None or len(self.results) < max_evals:
            path = self._model_dir_generator(self._next_step_to_fetch)
            mtime = os.path.getmtime(path) if os.path.isfile(path) else None
            if mtime is not None and mtime != self._last_mtime:
                self._last_mtime = mtime
                try:
                    self._load_model(path)
                except Exception as e:      # torn read of a non-atomic writer: retry
                    print("evaluator: could not load {} ({}), retrying".format(path, e))
                    time.sleep(min(1.0, self.poll_s))
                    continue
                print("Evaluator evaluating results on step {}".format(self._next_step_to_fetch))
                res = self._evaluate_model(validation_loader)
                res["step"] = self._next_step_to_fetch
                self.results.append(res)
                self._next_step_to_fetch += self._eval_freq
            else:
                if deadline is not None and time.time() > deadline:
                    break
                time.sleep(self.poll_s)
        return self.results
