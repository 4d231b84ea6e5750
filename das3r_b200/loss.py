"""Fused DAS3R photometric loss: train_gui.py:565-577 (static-mask multiply, L1, SSIM with the 11x11 Gaussian window,
PSNR, mean) as one autograd function over two forward kernels and one backward kernel (include/das3r_loss.h,
das3r_b200/csrc/loss.cu) instead of about forty torch kernels and their autograd graph.

    loss, psnr = photometric_loss(image, gt_image, static, lambda_dssim=opt.lambda_dssim)
    loss.backward()        # gradients reach `image` and `static` (and `gt_image` if it requires grad)

`static` is `gaussians._conf_static[viewpoint_cam.uid]` ([H,W]) or None. `psnr` is `psnr(image, gt).mean()` of the
masked images (train_gui.py:573), not differentiable.
"""
from __future__ import annotations

import torch

from . import load_native

_C = load_native()


class _PhotometricLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, gt, static, lambda_dssim):
        image_c, gt_c = image.contiguous(), gt.contiguous()
        static_c = static.contiguous() if static is not None else torch.empty(0, device=image.device)
        out, state = _C.loss_forward(image_c, gt_c, static_c, float(lambda_dssim))
        ctx.save_for_backward(image_c, gt_c, static_c, state)
        ctx.lambda_dssim = float(lambda_dssim)
        ctx.static_shape = None if static is None else static.shape
        psnr = out[1]
        ctx.mark_non_differentiable(psnr)
        return out[0], psnr

    @staticmethod
    def backward(ctx, grad_loss, _grad_psnr):
        image, gt, static, state = ctx.saved_tensors
        want_static = ctx.static_shape is not None and ctx.needs_input_grad[2]
        want_gt = ctx.needs_input_grad[1]
        g = grad_loss.reshape(1).to(torch.float32)
        dimg, dstatic, dgt = _C.loss_backward(image, gt, static, ctx.lambda_dssim, state, g, want_static, want_gt)
        return (dimg if ctx.needs_input_grad[0] else None, dgt if want_gt else None,
                dstatic.view(ctx.static_shape) if want_static else None, None)


def photometric_loss(image: torch.Tensor, gt: torch.Tensor, static: torch.Tensor | None = None,
                     lambda_dssim: float = 0.2):
    """-> (loss, psnr). image, gt: [3,H,W] float32 CUDA tensors; static: [H,W] (or [1,H,W]) or None."""
    if not image.is_cuda:
        raise RuntimeError("image must be a CUDA tensor: the loss has no CPU path")
    if image.dim() != 3 or image.shape[0] != 3 or gt.shape != image.shape:
        raise ValueError("image and gt must both be [3,H,W]")
    if static is not None and static.numel() != image.shape[1] * image.shape[2]:
        raise ValueError("static must have one value per pixel ([H,W])")
    return _PhotometricLoss.apply(image, gt, static, lambda_dssim)
