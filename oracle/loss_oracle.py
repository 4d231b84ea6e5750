"""CPU restatement (numpy) of DAS3R's photometric loss and its gradient - TEST INFRASTRUCTURE, like everything under
oracle/: only tests/, __graft_entry__.smoke() and bench.py's baseline legs may import it; das3r_b200/ never does.

Restated, with the reference lines (paths relative to /root/reference):
  image*static, gt*static                              train_gui.py:567-569
  Ll1 = |image - gt| (no reduction)                    utils/loss_utils.py:17-21
  window = outer(g, g), g = gaussian(11, 1.5)          utils/loss_utils.py:26-34
  ssim map (zero-padded depthwise conv, C1, C2)        utils/loss_utils.py:36-67
  psnr = mean_c 20 log10(1/sqrt(mse_c))                utils/image_utils.py:17-19, train_gui.py:573
  loss = mean((1-l)*Ll1 + l*(1-Lssim))                 train_gui.py:575-576
The gradient is the analytic adjoint (the reference gets it from autograd).

Parity pinned: tests/golden/loss_*.npz hold the outputs and autograd gradients of the reference's own
utils/loss_utils.py and utils/image_utils.py (imported from /root/reference by tests/golden/make_loss_golden.py, CPU
fp32); tests/test_loss_oracle.py checks this file against them.
"""
from __future__ import annotations

from math import exp

import numpy as np

C1, C2 = 0.01 ** 2, 0.03 ** 2


def gaussian_window(dtype=np.float64) -> np.ndarray:
    """utils/loss_utils.py:26-28: 1-D weights (float32 tensor divided by its float32 sum)."""
    g = np.array([exp(-(x - 11 // 2) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)], dtype=np.float32)
    return (g / g.sum(dtype=np.float32)).astype(dtype)


def _conv(x: np.ndarray, w2d: np.ndarray) -> np.ndarray:
    """Depthwise 11x11 correlation with zero padding 5 (F.conv2d(..., padding=5, groups=C)); x [C,H,W]."""
    C, H, W = x.shape
    xp = np.zeros((C, H + 10, W + 10), x.dtype)
    xp[:, 5:5 + H, 5:5 + W] = x
    out = np.zeros_like(x)
    for dy in range(11):
        for dx in range(11):
            out += w2d[dy, dx] * xp[:, dy:dy + H, dx:dx + W]
    return out


def forward(image, gt, static=None, lambda_dssim=0.2, dtype=np.float64) -> dict:
    f = lambda a: np.asarray(a, dtype=dtype)
    s = f(static).reshape(1, *np.shape(image)[1:]) if static is not None else dtype(1)
    x1, x2 = f(image) * s, f(gt) * s
    g = gaussian_window(dtype)
    w = np.outer(g, g)
    mu1, mu2 = _conv(x1, w), _conv(x2, w)
    e11, e22, e12 = _conv(x1 * x1, w), _conv(x2 * x2, w), _conv(x1 * x2, w)
    s1, s2, s12 = e11 - mu1 * mu1, e22 - mu2 * mu2, e12 - mu1 * mu2
    A1, A2 = 2 * mu1 * mu2 + C1, 2 * s12 + C2
    B1, B2 = mu1 * mu1 + mu2 * mu2 + C1, s1 + s2 + C2
    S = (A1 * A2) / (B1 * B2)
    l1 = np.abs(x1 - x2)
    loss = ((1.0 - lambda_dssim) * l1 + lambda_dssim * (1.0 - S)).mean()
    mse = ((x1 - x2) ** 2).reshape(x1.shape[0], -1).mean(1)
    psnr = (20 * np.log10(1.0 / np.sqrt(mse))).mean()
    return dict(loss=loss, psnr=psnr, ssim_map=S, x1=x1, x2=x2, w=w, mu1=mu1, mu2=mu2, A1=A1, A2=A2, B1=B1, B2=B2)


def backward(image, gt, static=None, lambda_dssim=0.2, grad_loss=1.0, dtype=np.float64) -> dict:
    f = lambda a: np.asarray(a, dtype=dtype)
    fw = forward(image, gt, static, lambda_dssim, dtype)
    x1, x2, w, mu1, mu2 = fw["x1"], fw["x2"], fw["w"], fw["mu1"], fw["mu2"]
    A1, A2, B1, B2, S = fw["A1"], fw["A2"], fw["B1"], fw["B2"], fw["ssim_map"]
    n = x1.size
    gS = -lambda_dssim * grad_loss / n
    gL = (1.0 - lambda_dssim) * grad_loss / n
    inv = 1.0 / (B1 * B2)
    common = 2 * (A2 - A1) * inv
    d_mu1 = mu2 * common - 2 * mu1 * S / B1 + 2 * mu1 * S / B2
    d_mu2 = mu1 * common - 2 * mu2 * S / B1 + 2 * mu2 * S / B2
    d_e11 = -S / B2
    d_e22 = -S / B2
    d_e12 = 2 * A1 * inv
    sg = np.sign(x1 - x2)
    # the window is symmetric, so the adjoint of the zero-padded correlation is the same correlation
    gx1 = gS * (_conv(d_mu1, w) + 2 * x1 * _conv(d_e11, w) + x2 * _conv(d_e12, w)) + gL * sg
    gx2 = gS * (_conv(d_mu2, w) + 2 * x2 * _conv(d_e22, w) + x1 * _conv(d_e12, w)) - gL * sg
    s = f(static).reshape(1, *x1.shape[1:]) if static is not None else np.ones((1, *x1.shape[1:]), dtype)
    return dict(dL_dimage=gx1 * s, dL_dgt=gx2 * s, dL_dstatic=(gx1 * f(image) + gx2 * f(gt)).sum(0))
