"""CPU restatement (numpy) of one torch.optim.Adam step as DAS3R configures it (scene/gaussian_model.py:258:
`torch.optim.Adam(l, lr=0.0, eps=1e-15)`, per-group learning rates, no weight decay / amsgrad) - TEST INFRASTRUCTURE
(see oracle/__init__.py). Follows torch/optim/adam.py `_single_tensor_adam`. Pinned by tests/test_adam_oracle.py against
torch.optim.Adam itself (the reference's optimizer IS that library class; it is importable on the CPU)."""
from __future__ import annotations

import numpy as np


def adam_step(p, g, m, v, lr, step, beta1=0.9, beta2=0.999, eps=1e-15, dtype=np.float64):
    """-> (p, m, v) after the update; `step` is the 1-based count of this update."""
    p, g, m, v = (np.asarray(a, dtype=dtype) for a in (p, g, m, v))
    m = m + (1 - beta1) * (g - m)                      # exp_avg.lerp_(grad, 1 - beta1)
    v = v * beta2 + (1 - beta2) * g * g                # exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
    bc1, bc2 = 1 - beta1 ** step, 1 - beta2 ** step
    denom = np.sqrt(v) / np.sqrt(bc2) + eps
    return p - (lr / bc1) * (m / denom), m, v
