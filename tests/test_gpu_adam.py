"""FusedAdam (include/das3r_adam.h) on the GPU against torch.optim.Adam configured as DAS3R does
(scene/gaussian_model.py:236-258) and against the float64 oracle, over several steps with changing learning rates,
missing gradients, unaligned / odd-sized tensors and more tensors than one launch takes."""
import numpy as np
import pytest
import torch

from oracle import adam_oracle as ao

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
LRS = [1.6e-4, 2.5e-3, 2.5e-3 / 20, 0.05, 5e-3, 1e-3, 3e-3]
SHAPES = [(10_007, 3), (10_007, 1, 3), (10_007, 15, 3), (10_007, 1), (10_007, 3), (10_007, 4), (5, 36, 64)]


def test_matches_torch_adam_and_the_oracle():
    from das3r_b200.optim import FusedAdam

    g = torch.Generator().manual_seed(1)
    init = [torch.randn(*s, generator=g) for s in SHAPES]
    a = [t.clone().to(DEV).requires_grad_(True) for t in init]
    b = [t.clone().to(DEV).requires_grad_(True) for t in init]
    opt_a = torch.optim.Adam([{"params": [p], "lr": lr, "name": str(i)} for i, (p, lr) in enumerate(zip(a, LRS))], lr=0.0,
                             eps=1e-15)
    opt_b = FusedAdam([{"params": [p], "lr": lr, "name": str(i)} for i, (p, lr) in enumerate(zip(b, LRS))], lr=0.0,
                      eps=1e-15)
    mine = [(t.double().numpy().copy(), np.zeros(t.shape), np.zeros(t.shape), 0) for t in init]
    for step in range(1, 7):
        grads = [torch.randn(*s, generator=g) * (10.0 ** -(i % 4)) for i, s in enumerate(SHAPES)]
        skip = {3} if step == 2 else set()                 # a parameter without gradient is left alone, step not counted
        for i, (pa, pb, gr) in enumerate(zip(a, b, grads)):
            pa.grad = None if i in skip else gr.to(DEV)
            pb.grad = None if i in skip else gr.to(DEV)
        if step == 4:                                      # the schedulers rewrite param_group['lr'] (gaussian_model.py:276-310)
            for og in (opt_a, opt_b):
                og.param_groups[0]["lr"] = 7e-5
        opt_a.step()
        opt_b.step()
        lrs = [opt_b.param_groups[i]["lr"] for i in range(len(b))]
        new = []
        for i, ((p, m, v, t), gr) in enumerate(zip(mine, grads)):
            if i in skip:
                new.append((p, m, v, t))
            else:
                p2, m2, v2 = ao.adam_step(p, gr.numpy(), m, v, lrs[i], t + 1)
                new.append((p2, m2, v2, t + 1))
        mine = new
        torch.cuda.synchronize()
        for i, (pa, pb) in enumerate(zip(a, b)):
            np.testing.assert_allclose(pb.detach().cpu().numpy(), pa.detach().cpu().numpy(), rtol=2e-6, atol=2e-7)
            np.testing.assert_allclose(pb.detach().cpu().numpy(), mine[i][0], rtol=1e-5, atol=1e-6)
            if i not in skip or step > 2:
                sa, sb = opt_a.state[pa], opt_b.state[pb]
                assert float(sa["step"]) == float(sb["step"])
                np.testing.assert_allclose(sb["exp_avg"].cpu().numpy(), sa["exp_avg"].cpu().numpy(), rtol=2e-6, atol=1e-9)
                np.testing.assert_allclose(sb["exp_avg_sq"].cpu().numpy(), sa["exp_avg_sq"].cpu().numpy(), rtol=2e-6,
                                           atol=1e-12)


def test_more_tensors_than_one_launch_and_unaligned_views():
    from das3r_b200.optim import FusedAdam

    g = torch.Generator().manual_seed(2)
    base = torch.randn(40, 1001, generator=g)
    storage_a, storage_b = base.clone().to(DEV), base.clone().to(DEV)
    # 20 parameters that are rows of one buffer starting at odd element offsets: not 16-byte aligned
    a = [storage_a[i, 1:].detach().requires_grad_(True) for i in range(20)]
    b = [storage_b[i, 1:].detach().requires_grad_(True) for i in range(20)]
    opt_a = torch.optim.Adam(a, lr=1e-2, eps=1e-15)
    opt_b = FusedAdam(b, lr=1e-2, eps=1e-15)
    for _ in range(3):
        for pa, pb in zip(a, b):
            gr = torch.randn(1000, generator=g).to(DEV)
            pa.grad, pb.grad = gr, gr.clone()
        opt_a.step()
        opt_b.step()
    torch.cuda.synchronize()
    for pa, pb in zip(a, b):
        np.testing.assert_allclose(pb.detach().cpu().numpy(), pa.detach().cpu().numpy(), rtol=2e-6, atol=2e-7)
    assert torch.equal(storage_a[:, 0], storage_b[:, 0])  # the elements between the views were not touched


def test_rejects_cpu_parameters():
    from das3r_b200.optim import FusedAdam

    p = torch.zeros(4, requires_grad=True)
    p.grad = torch.ones(4)
    with pytest.raises(RuntimeError):
        FusedAdam([p], lr=1e-3).step()
