"""``from nn_ops import NN_Trainer`` (src/nn_ops.py)."""
from ewdml.train.roles import NNTrainer as _T
from ewdml.utils.metrics import accuracy  # noqa: F401
from _cfg import config_from_kwargs


class NN_Trainer(_T):
    def __init__(self, **kwargs):
        super().__init__(config_from_kwargs(kwargs), kwargs.get("device"))
