"""Generates tests/golden/loss_*.npz: loss, psnr, SSIM map and autograd gradients computed by the REFERENCE's own
utils/loss_utils.py (l1_loss, ssim) and utils/image_utils.py (psnr), imported from /root/reference as they are, glued
exactly like train_gui.py:565-576 (CPU, fp32). Run where /root/reference exists:
python tests/golden/make_loss_golden.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from utils.loss_utils import l1_loss, ssim  # noqa: E402
from utils.image_utils import psnr  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def make(name, H, W, seed, lam, with_static=True, grad_scale=1.0):
    g = torch.Generator().manual_seed(seed)
    # smooth-ish images in [0,1] plus noise, so that SSIM is neither 0 nor 1 everywhere
    yy, xx = torch.meshgrid(torch.linspace(0, 3, H), torch.linspace(0, 4, W), indexing="ij")
    base = torch.stack([0.5 + 0.4 * torch.sin(xx + c) * torch.cos(yy * (1 + 0.3 * c)) for c in range(3)])
    gt0 = (base + 0.05 * torch.randn(3, H, W, generator=g)).clamp(0, 1)
    img0 = (base + 0.1 * torch.randn(3, H, W, generator=g)).clamp(0, 1)
    img0[:, : H // 3] = gt0[:, : H // 3]  # a region where image == gt exactly (|x| at 0, SSIM = 1)
    static0 = (0.3 + 0.7 * torch.rand(H, W, generator=g)) if with_static else None
    image = img0.clone().requires_grad_(True)
    gt_image = gt0.clone().requires_grad_(True)
    static = static0.clone().requires_grad_(True) if with_static else None
    im, gti = (image * static, gt_image * static) if with_static else (image, gt_image)   # train_gui.py:568-569
    Ll1 = l1_loss(im, gti, reduce=False)                                                  # :570
    Lssim = ssim(im, gti, size_average=False)                                             # :571
    psnr_frame = psnr(im, gti).mean()                                                     # :573
    loss = ((1.0 - lam) * Ll1 + lam * (1.0 - Lssim)).mean()                               # :575-576
    (loss * grad_scale).backward()
    save = dict(image=img0.numpy(), gt=gt0.numpy(), lambda_dssim=np.float32(lam), grad_scale=np.float32(grad_scale),
                loss=loss.detach().numpy(), psnr=psnr_frame.detach().numpy(), ssim_map=Lssim.detach().numpy(),
                grad_image=image.grad.numpy(), grad_gt=gt_image.grad.numpy())
    if with_static:
        save.update(static=static0.numpy(), grad_static=static.grad.numpy())
    np.savez_compressed(os.path.join(HERE, f"loss_{name}.npz"), **save)
    print(name, "loss", float(loss), "psnr", float(psnr_frame))


if __name__ == "__main__":
    make("small", H=40, W=56, seed=0, lam=0.2)
    make("odd_nostatic", H=37, W=51, seed=1, lam=0.2, with_static=False)
    make("tiny", H=7, W=9, seed=2, lam=0.5, grad_scale=3.0)     # smaller than the window: padding everywhere
    make("frame", H=96, W=160, seed=3, lam=0.2)
