"""oracle/adam_oracle.py against torch.optim.Adam configured as in scene/gaussian_model.py:236-258 (seven groups, one
learning rate each, eps 1e-15), over several steps, on the CPU."""
import numpy as np
import torch

from oracle import adam_oracle as ao

LRS = [1.6e-4, 2.5e-3, 2.5e-3 / 20, 0.05, 5e-3, 1e-3, 3e-3]          # arguments/__init__.py:72-83 + conf_lr_init
SHAPES = [(50, 3), (50, 1, 3), (50, 15, 3), (50, 1), (50, 3), (50, 4), (4, 6, 8)]


def test_matches_torch_adam_over_steps():
    g = torch.Generator().manual_seed(0)
    params = [torch.randn(*s, generator=g).requires_grad_(True) for s in SHAPES]
    opt = torch.optim.Adam([{"params": [p], "lr": lr} for p, lr in zip(params, LRS)], lr=0.0, eps=1e-15)
    mine = [(p.detach().double().numpy().copy(), np.zeros(p.shape), np.zeros(p.shape)) for p in params]
    for step in range(1, 6):
        grads = [torch.randn(*s, generator=g) * (10.0 ** -(i % 4)) for i, s in enumerate(SHAPES)]
        for p, gr in zip(params, grads):
            p.grad = gr.clone()
        opt.step()
        mine = [ao.adam_step(p, gr.numpy(), m, v, lr, step) for (p, m, v), gr, lr in zip(mine, grads, LRS)]
        for p, (q, m, v) in zip(params, mine):
            np.testing.assert_allclose(p.detach().numpy(), q, rtol=1e-5, atol=1e-7)
            np.testing.assert_allclose(opt.state[p]["exp_avg"].numpy(), m, rtol=1e-5, atol=1e-8)
            np.testing.assert_allclose(opt.state[p]["exp_avg_sq"].numpy(), v, rtol=1e-5, atol=1e-11)
