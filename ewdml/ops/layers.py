"""Layer modules used by every model in :mod:`ewdml.models`.

They subclass the ``torch.nn`` modules of the same name, so parameter names, shapes and
``state_dict`` keys are those of the reference models (``src/model_ops/*.py``) and checkpoints
load both ways.  What differs is *execution*: when a module's ``compute`` is ``"native"`` and the
input lives on a CUDA device, forward/backward run the hand-written sm_100a kernels in
:mod:`ewdml.ops.native` (tcgen05 GEMM, implicit-GEMM convolution, fused BatchNorm/ReLU, pooling);
otherwise they defer to torch (CPU plumbing path and the eager baseline).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

_COMPUTE = {"mode": "torch"}


def set_compute(mode: str) -> None:
    """Select ``"torch"`` or ``"native"`` execution for all ewdml layers (process-wide)."""
    if mode not in ("torch", "native"):
        raise ValueError(mode)
    _COMPUTE["mode"] = mode


def get_compute() -> str:
    return _COMPUTE["mode"]


def _native(x: torch.Tensor) -> bool:
    return _COMPUTE["mode"] == "native" and x.is_cuda


class Conv2d(nn.Conv2d):
    pass


class Linear(nn.Linear):
    def forward(self, x):
        if _native(x):
            from . import native
            return native.linear_block(x, self, relu=False)
        return super().forward(x)


class BatchNorm2d(nn.BatchNorm2d):
    pass


class ReLU(nn.ReLU):
    def forward(self, x):
        if _native(x) and not self.inplace:
            from . import native
            return native.relu(x)
        return super().forward(x)


class MaxPool2d(nn.MaxPool2d):
    def forward(self, x):
        k = self.kernel_size if isinstance(self.kernel_size, int) else self.kernel_size[0]
        st = self.stride if isinstance(self.stride, int) else self.stride[0]
        if _native(x) and x.dim() == 4 and k == 2 and st == 2 and not self.padding and not self.ceil_mode:
            from . import native
            return native.max_pool2x2(x)
        return super().forward(x)


class AvgPool2d(nn.AvgPool2d):
    def forward(self, x):
        if _native(x):
            from . import native
            return native.avg_pool(x, self.kernel_size)
        return super().forward(x)


class Dropout(nn.Dropout):
    def forward(self, x):
        if _native(x) and self.training and self.p > 0:
            from . import native
            return native.dropout(x, self.p)
        return super().forward(x)


def conv_bn_act(x, conv: nn.Conv2d, bn=None, relu: bool = False, residual=None, pool: bool = False):
    """``[maxpool2x2] relu?( bn?( conv(x) ) [+ residual] )`` -- the unit every model here is built from.

    Native mode: one fused chain of sm_100a kernels (im2col -> tcgen05 GEMM -> BN statistics ->
    normalise/add/activate/pool).  Torch mode: the same modules called one after another.
    """
    if _native(x):
        from . import native
        return native.conv_block(x, conv, bn, relu=relu, residual=residual, pool=pool)
    y = conv(x)
    if bn is not None:
        y = bn(y)
    if residual is not None:
        y = y + residual
    if relu:
        y = F.relu(y)
    if pool:
        y = F.max_pool2d(y, 2, 2)
    return y


def linear_act(x, lin: nn.Linear, relu: bool = False):
    if _native(x):
        from . import native
        return native.linear_block(x, lin, relu=relu)
    y = lin(x)
    return F.relu(y) if relu else y


def cross_entropy(logits: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """Mean softmax cross-entropy (``nn.CrossEntropyLoss`` at ``distributed_worker.py:103``)."""
    if _native(logits):
        from . import native
        return native.cross_entropy(logits, target)
    return F.cross_entropy(logits.float(), target)
