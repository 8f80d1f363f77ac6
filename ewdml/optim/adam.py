"""Adam taking external gradients (``src/optim/adam.py:12-95``).

The reference expects numpy gradients (``adam.py:50``); tensors and numpy arrays are both accepted
here.  Standard bias-corrected Adam with optional AMSGrad and L2 weight decay.
"""
from __future__ import annotations

import math
from typing import Optional, Sequence

import torch
from torch.optim import Optimizer


class Adam(Optimizer):
    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, amsgrad: bool = False):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                                      amsgrad=amsgrad))

    @torch.no_grad()
    def step(self, grads: Optional[Sequence] = None, closure=None):
        loss = closure() if closure is not None else None
        offset = 0
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for i, p in enumerate(group["params"]):
                g = grads[offset + i] if grads is not None else p.grad
                if g is None:
                    continue
                g = torch.as_tensor(g).to(device=p.device, dtype=p.dtype).view_as(p)
                st = self.state[p]
                if not st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p)
                    st["exp_avg_sq"] = torch.zeros_like(p)
                    if group["amsgrad"]:
                        st["max_exp_avg_sq"] = torch.zeros_like(p)
                st["step"] += 1
                if group["weight_decay"] != 0:
                    g = g.add(p, alpha=group["weight_decay"])
                st["exp_avg"].mul_(b1).add_(g, alpha=1 - b1)
                st["exp_avg_sq"].mul_(b2).addcmul_(g, g, value=1 - b2)
                if group["amsgrad"]:
                    torch.maximum(st["max_exp_avg_sq"], st["exp_avg_sq"], out=st["max_exp_avg_sq"])
                    denom = st["max_exp_avg_sq"].sqrt().add_(group["eps"])
                else:
                    denom = st["exp_avg_sq"].sqrt().add_(group["eps"])
                bc1 = 1 - b1 ** st["step"]
                bc2 = 1 - b2 ** st["step"]
                p.addcdiv_(st["exp_avg"], denom, value=-group["lr"] * math.sqrt(bc2) / bc1)
            offset += len(group["params"])
        return loss


class FlatAdam:
    """Adam on the flat parameter buffer of :class:`ewdml.parallel.flat.FlatModel` (one fused sequence of whole-model
    tensor ops instead of a Python loop over parameters).  The step counter lives in a device tensor and every operation
    is a tensor op, so the update can be captured in a CUDA graph together with the exchange that produced the gradient.
    Same arithmetic as :class:`Adam` (bias-corrected, optional L2 weight decay); used by the native exchanges when
    ``--optimizer adam`` is selected (their fused epilogue implements momentum-SGD only)."""

    def __init__(self, param: torch.Tensor, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, param_bf16: Optional[torch.Tensor] = None):
        self.param, self.param_bf16 = param, param_bf16
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.exp_avg = torch.zeros_like(param)
        self.exp_avg_sq = torch.zeros_like(param)
        self.t = torch.zeros((), dtype=torch.float64, device=param.device)

    @torch.no_grad()
    def step(self, grad: torch.Tensor) -> None:
        b1, b2 = self.betas
        self.t.add_(1.0)
        g = grad if self.weight_decay == 0 else grad.add(self.param, alpha=self.weight_decay)
        self.exp_avg.mul_(b1).add_(g, alpha=1 - b1)
        self.exp_avg_sq.mul_(b2).addcmul_(g, g, value=1 - b2)
        bc1 = 1.0 - torch.pow(float(b1), self.t)          # scalar base ** device exponent: no host -> device copy in the step
        bc2 = 1.0 - torch.pow(float(b2), self.t)
        step_size = (self.lr * torch.sqrt(bc2) / bc1).to(torch.float32)
        self.param.sub_(self.exp_avg / self.exp_avg_sq.sqrt().add_(self.eps) * step_size)
        if self.param_bf16 is not None:
            self.param_bf16.copy_(self.param)

    def state_dict(self):
        return {"exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq, "t": self.t}

    def load_state_dict(self, sd) -> None:
        self.exp_avg.copy_(sd["exp_avg"]); self.exp_avg_sq.copy_(sd["exp_avg_sq"]); self.t.copy_(sd["t"])
