"""``from distributed_worker import *`` (src/distributed_worker.py)."""
import torch

from ewdml.train.roles import DistributedWorker as _Worker
from ewdml.utils.metrics import accuracy  # noqa: F401
from _cfg import config_from_kwargs


class ModelBuffer(object):
    """Receive buffers for the model pulled from the server (``distributed_worker.py:41-59``): here one flat allocation,
    ``recv_buf[i]`` being a view shaped like parameter ``i``."""

    def __init__(self, network):
        params = list(network.parameters())
        self._flat = torch.zeros(sum(p.numel() for p in params))
        self.recv_buf, self.layer_cur_step, self.layer_shape, off = [], [], [], 0
        for p in params:
            self.recv_buf.append(self._flat[off:off + p.numel()].view(p.shape))
            self.layer_cur_step.append(0)
            self.layer_shape.append(p.shape)
            off += p.numel()


class DistributedWorker(_Worker):
    def __init__(self, **kwargs):
        super().__init__(config_from_kwargs(kwargs), kwargs["rank"], kwargs["world_size"], kwargs.get("device"))

    def train_updated(self, train_loader, test_loader=None):
        return self.train(train_loader, test_loader)
