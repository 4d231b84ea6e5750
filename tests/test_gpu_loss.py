"""The fused photometric loss (include/das3r_loss.h) on the GPU against the golden vectors of the reference's own loss
code (tests/golden/loss_*.npz), the float64 oracle at frame size, and the reference formulation in torch on the GPU
through autograd. Tolerances: loss 1e-5 absolute (a mean of O(0.1) values), psnr 1e-3 dB, gradients 1e-3 relative to the
tensor's largest entry - the reference itself forms E[x^2] - mu^2 in fp32, which leaves ~1e-4 noise in single pixels of
the SSIM map."""
import glob
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import _harness as Hn
from oracle import loss_oracle as lo

pytestmark = pytest.mark.gpu
GOLDEN = sorted(glob.glob(os.path.join(Hn.ROOT, "tests", "golden", "loss_*.npz")))
DEV = "cuda:0"


def _t(a):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32, device=DEV)


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_against_reference_golden(path):
    from das3r_b200.loss import photometric_loss

    z = np.load(path)
    image, gt = _t(z["image"]).requires_grad_(True), _t(z["gt"]).requires_grad_(True)
    static = _t(z["static"]).requires_grad_(True) if "static" in z.files else None
    loss, psnr = photometric_loss(image, gt, static, float(z["lambda_dssim"]))
    (loss * float(z["grad_scale"])).backward()
    torch.cuda.synchronize()
    assert abs(float(loss.detach()) - float(z["loss"])) <= 1e-5
    assert abs(float(psnr) - float(z["psnr"])) <= 1e-3
    assert Hn.rel_err(image.grad, z["grad_image"]) <= 1e-3
    assert Hn.rel_err(gt.grad, z["grad_gt"]) <= 1e-3
    if static is not None:
        assert Hn.rel_err(static.grad, z["grad_static"]) <= 1e-3


def _frame(H, W, seed):
    r = np.random.default_rng(seed)
    yy, xx = np.meshgrid(np.linspace(0, 3, H), np.linspace(0, 4, W), indexing="ij")
    base = np.stack([0.5 + 0.4 * np.sin(xx + c) * np.cos(yy * (1 + 0.3 * c)) for c in range(3)])
    gt = np.clip(base + 0.05 * r.standard_normal((3, H, W)), 0, 1).astype(np.float32)
    img = np.clip(base + 0.1 * r.standard_normal((3, H, W)), 0, 1).astype(np.float32)
    return img, gt, (0.3 + 0.7 * r.random((H, W))).astype(np.float32)


@pytest.mark.parametrize("H,W", [(288, 512), (33, 17), (1, 1)])
def test_against_the_oracle(H, W):
    from das3r_b200.loss import photometric_loss

    img, gt, static = _frame(H, W, seed=H + W)
    image, g, s = _t(img).requires_grad_(True), _t(gt).requires_grad_(True), _t(static).requires_grad_(True)
    loss, psnr = photometric_loss(image, g, s, 0.2)
    loss.backward()
    torch.cuda.synchronize()
    fw = lo.forward(img, gt, static, 0.2)
    bw = lo.backward(img, gt, static, 0.2)
    assert abs(float(loss.detach()) - fw["loss"]) <= 1e-5
    assert abs(float(psnr) - fw["psnr"]) <= 1e-3
    assert Hn.rel_err(image.grad, bw["dL_dimage"]) <= 1e-3
    assert Hn.rel_err(g.grad, bw["dL_dgt"]) <= 1e-3
    assert Hn.rel_err(s.grad, bw["dL_dstatic"]) <= 1e-3


def test_matches_the_reference_formulation_in_torch_and_is_deterministic():
    from das3r_b200.loss import photometric_loss

    H, W = 288, 512
    img, gt, static = _frame(H, W, seed=7)

    def torch_loss(image, gt_image, st, lam=0.2):  # train_gui.py:567-576 with utils/loss_utils.py:36-67
        im, g = image * st, gt_image * st
        w1 = torch.tensor(lo.gaussian_window(np.float32), device=DEV)
        window = (w1[:, None] @ w1[None, :])[None, None].expand(3, 1, 11, 11).contiguous()
        conv = lambda x: F.conv2d(x, window, padding=5, groups=3)
        mu1, mu2 = conv(im), conv(g)
        s1, s2, s12 = conv(im * im) - mu1.pow(2), conv(g * g) - mu2.pow(2), conv(im * g) - mu1 * mu2
        C1, C2 = 0.01 ** 2, 0.03 ** 2
        ssim_map = ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1.pow(2) + mu2.pow(2) + C1) * (s1 + s2 + C2))
        return ((1.0 - lam) * torch.abs(im - g) + lam * (1.0 - ssim_map)).mean()

    a = [_t(x).requires_grad_(True) for x in (img, gt, static)]
    torch_loss(*a).backward()
    b = [_t(x).requires_grad_(True) for x in (img, gt, static)]
    loss, _ = photometric_loss(*b, 0.2)
    loss.backward()
    c = [_t(x).requires_grad_(True) for x in (img, gt, static)]
    loss2, _ = photometric_loss(*c, 0.2)
    loss2.backward()
    torch.cuda.synchronize()
    assert abs(float(loss.detach()) - float(torch_loss(*[t.detach() for t in a]))) <= 1e-5
    for x, y in zip(b, a):
        assert Hn.rel_err(x.grad, y.grad) <= 1e-3
    assert float(loss.detach()) == float(loss2.detach())  # fixed-order reduction: run-to-run identical
    for x, y in zip(b, c):
        assert torch.equal(x.grad, y.grad)
