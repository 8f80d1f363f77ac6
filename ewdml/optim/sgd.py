"""SGD that consumes externally supplied gradients: ``optimizer.step(grads=[...])``.

API and update rule of ``src/optim/sgd.py:59-92``: gradient ``i`` of the list belongs to parameter
``i`` of the group; L2 weight decay is added to the gradient; the momentum buffer is initialised to
the first gradient (no dampening on the first step) and afterwards ``buf = mu*buf + (1-damp)*g``;
Nesterov uses ``g + mu*buf``; then ``p -= lr * d``.  ``grads=None`` falls back to ``p.grad``.

On CUDA with a :class:`ewdml.parallel.flat.FlatModel` the same rule runs as one fused sm_100a
kernel over the flat parameter/momentum buffers (``csrc/comm/apply.cu``); this class is the
reference semantics and the CPU path.
"""
from __future__ import annotations

from typing import Iterable, Optional, Sequence

import torch
from torch.optim import Optimizer


class SGD(Optimizer):
    def __init__(self, params, lr: float = 0.1, momentum: float = 0.0, dampening: float = 0.0,
                 weight_decay: float = 0.0, nesterov: bool = False):
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        super().__init__(params, dict(lr=lr, momentum=momentum, dampening=dampening,
                                      weight_decay=weight_decay, nesterov=nesterov))

    @torch.no_grad()
    def step(self, grads: Optional[Sequence[torch.Tensor]] = None, closure=None):
        loss = closure() if closure is not None else None
        offset = 0
        for group in self.param_groups:
            mu, damp = group["momentum"], group["dampening"]
            wd, lr, nesterov = group["weight_decay"], group["lr"], group["nesterov"]
            for i, p in enumerate(group["params"]):
                g = grads[offset + i] if grads is not None else p.grad
                if g is None:
                    continue
                if isinstance(g, torch.Tensor) is False:
                    g = torch.as_tensor(g)
                g = g.to(device=p.device, dtype=p.dtype).view_as(p)
                if wd != 0:
                    g = g.add(p, alpha=wd)
                if mu != 0:
                    st = self.state[p]
                    buf = st.get("momentum_buffer")
                    if buf is None:
                        buf = st["momentum_buffer"] = g.clone()
                    else:
                        buf.mul_(mu).add_(g, alpha=1.0 - damp)
                    g = g.add(buf, alpha=mu) if nesterov else buf
                p.add_(g, alpha=-lr)
            offset += len(group["params"])
        return loss
