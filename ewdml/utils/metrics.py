"""Per-step metrics: accuracy, structured JSONL records, and the reference-compatible log line.

``accuracy`` is precision@k as in ``distributed_worker.py:27-39``.  ``StepLogger`` writes one JSON
object per step (step, loss, top-1/5, images/s, exact payload bytes sent/received, compression
ratio, per-phase device milliseconds) and a rank-0 human line carrying the reference's fields
(``distributed_worker.py:230-231``: worker, step, total receive bytes, total send bytes, total
receive time, total send time, accuracy).  Byte counts are exact wire bytes, not
``sys.getsizeof(storage)``.
"""
from __future__ import annotations

import json
import logging
import time
from typing import Dict, Optional, Sequence

import torch

logger = logging.getLogger("ewdml")
if not logger.handlers:
    _h = logging.StreamHandler()
    _h.setFormatter(logging.Formatter("%(levelname)s:%(name)s:%(message)s"))
    logger.addHandler(_h)
    logger.setLevel(logging.INFO)
    logger.propagate = False


@torch.no_grad()
def accuracy(output: torch.Tensor, target: torch.Tensor, topk: Sequence[int] = (1,)):
    """precision@k in percent for each k (list of 1-element tensors, like the reference)."""
    maxk = min(max(topk), output.shape[1])
    pred = output.topk(maxk, dim=1, largest=True, sorted=True).indices.t()
    hit = pred.eq(target.view(1, -1).expand_as(pred))
    n = target.shape[0]
    return [hit[:min(k, maxk)].reshape(-1).float().sum(0, keepdim=True).mul_(100.0 / n) for k in topk]


class StepLogger:
    def __init__(self, rank: int, path: str = "", log_interval: int = 1):
        self.rank = rank
        self.path = path
        self.log_interval = max(1, log_interval)
        self._f = open(path, "a") if path else None
        self.total_sent = 0
        self.total_recv = 0
        self.time_send = 0.0
        self.time_recv = 0.0

    def log(self, step: int, *, loss: float, prec1: float, prec5: float = float("nan"),
            bytes_sent: int = 0, bytes_recv: int = 0, dense_bytes: int = 0, t_send: float = 0.0,
            t_recv: float = 0.0, t_compute: float = 0.0, images: int = 0,
            extra: Optional[Dict] = None) -> Dict:
        self.total_sent += bytes_sent
        self.total_recv += bytes_recv
        self.time_send += t_send
        self.time_recv += t_recv
        t_total = t_send + t_recv + t_compute
        rec = {"rank": self.rank, "step": step, "loss": loss, "prec1": prec1, "prec5": prec5,
               "bytes_sent": bytes_sent, "bytes_recv": bytes_recv,
               "compression_ratio": (dense_bytes / bytes_sent) if bytes_sent else None,
               "ms_send": t_send * 1e3, "ms_recv": t_recv * 1e3, "ms_compute": t_compute * 1e3,
               "images_per_s": (images / t_total) if t_total > 0 else None,
               "exchange_GBps": ((bytes_sent + bytes_recv) / (t_send + t_recv) / 1e9)
               if (t_send + t_recv) > 0 else None, "time": time.time()}
        if extra:
            rec.update(extra)
        if self._f:
            self._f.write(json.dumps(rec) + "\n")
            self._f.flush()
        if step % self.log_interval == 0:
            logger.info("Worker: %d, Step: %d, ,  total recieve memoryt = %d , total send memory = %d , "
                        "total time recive = %.6f , total_time send = %.6f , accuracy = %.4f , loss = %.4f",
                        self.rank, step, self.total_recv, self.total_sent, self.time_recv,
                        self.time_send, prec1, loss)
        return rec

    def close(self):
        if self._f:
            self._f.close()
            self._f = None
