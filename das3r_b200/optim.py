"""Fused Adam for DAS3R's optimizers: `FusedAdam(l, lr=0.0, eps=1e-15)` is call-compatible with the
`torch.optim.Adam(l, lr=0.0, eps=1e-15)` of scene/gaussian_model.py:258-260 - same param_groups (the lr schedulers of
gaussian_model.py:276-310 write `param_group['lr']` as before), same per-parameter state keys (`step`, `exp_avg`,
`exp_avg_sq`, so the state surgery helpers of gaussian_model.py keep working), same `step()` / `zero_grad()` - but all
parameter tensors are updated by ONE kernel launch (include/das3r_adam.h) instead of five to six passes per group.
No weight decay, amsgrad or maximize (DAS3R uses none); float32 CUDA parameters only."""
from __future__ import annotations

import torch

from . import load_native

_C = load_native()


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        if lr < 0.0 or eps < 0.0 or not (0.0 <= betas[0] < 1.0) or not (0.0 <= betas[1] < 1.0):
            raise ValueError("invalid Adam hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        # tensors that share (betas, eps, step) go into one launch; DAS3R's groups all do
        buckets = {}
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["step"] += 1
                key = (b1, b2, group["eps"], int(st["step"]))
                lst = buckets.setdefault(key, ([], [], [], [], []))
                if not (p.is_contiguous() and st["exp_avg"].is_contiguous() and st["exp_avg_sq"].is_contiguous()):
                    raise RuntimeError("FusedAdam needs contiguous parameters and moment buffers")
                lst[0].append(p)
                lst[1].append(p.grad.contiguous())
                lst[2].append(st["exp_avg"])
                lst[3].append(st["exp_avg_sq"])
                lst[4].append(float(group["lr"]))
        for (b1, b2, eps, step), (ps, gs, ms, vs, lrs) in buckets.items():
            _C.adam_step(ps, gs, ms, vs, lrs, b1, b2, eps, step)
        return loss
