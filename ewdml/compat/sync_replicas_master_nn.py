"""``from sync_replicas_master_nn import *`` (src/sync_replicas_master_nn.py)."""
import torch

from ewdml.train.roles import SyncReplicasMaster_NN as _Master
from ewdml.utils.metrics import accuracy  # noqa: F401
from _cfg import config_from_kwargs


def update_params_dist_version(param, avg_grad, learning_rate):
    """Plain SGD on one layer (``sync_replicas_master_nn.py:24-30``)."""
    if tuple(param.shape) != tuple(avg_grad.shape):
        raise ValueError(f"shape mismatch: {tuple(param.shape)} vs {tuple(avg_grad.shape)}")
    return param.sub_(avg_grad, alpha=float(learning_rate))


class GradientAccumulator(object):
    """Per-parameter, per-worker gradient slots (``sync_replicas_master_nn.py:49-86``); one [world][total] allocation,
    ``gradient_aggregator[i][w]`` is worker ``w``'s view of parameter ``i``."""

    def __init__(self, module, world_size, mode="None"):
        params = list(module.parameters())
        total = sum(p.numel() for p in params)
        self._flat = torch.zeros(world_size, total)
        self._mode = mode
        self.gradient_aggregator, self.gradient_aggregate_counter, self.model_index_range, off = [], [], [], 0
        for i, p in enumerate(params):
            self.gradient_aggregator.append([self._flat[w, off:off + p.numel()].view(p.shape) for w in range(world_size)])
            self.gradient_aggregate_counter.append(0)
            self.model_index_range.append(i)
            off += p.numel()

    # (method names, including the reference's spelling, kept for callers)
    def _meset_grad_counter(self):
        self.gradient_aggregate_counter = [0] * len(self.gradient_aggregator)

    def _meset_grad_aggregator(self):
        if self._mode != "compress":          # one memset of the single backing allocation clears every slot
            self._flat.zero_()

    def meset_everything(self):
        self._meset_grad_aggregator()
        self._meset_grad_counter()


class SyncReplicasMaster_NN(_Master):
    def __init__(self, **kwargs):
        super().__init__(config_from_kwargs(kwargs), kwargs["world_size"], kwargs.get("device"))

    def start_updated(self):
        return self.start()
