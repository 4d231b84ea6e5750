"""Developer tool (GPU box): fused photometric loss (forward + backward) against the reference formulation in torch
(train_gui.py:567-576 with utils/loss_utils.py) on one H x W frame. Usage: python tools/loss_bench.py [H] [W] [iters]"""
import json
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from das3r_b200.loss import photometric_loss  # noqa: E402


def main():
    H = int(sys.argv[1]) if len(sys.argv) > 1 else 288
    W = int(sys.argv[2]) if len(sys.argv) > 2 else 512
    iters = int(sys.argv[3]) if len(sys.argv) > 3 else 50
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    img, gt = torch.rand(3, H, W, device=dev, generator=g), torch.rand(3, H, W, device=dev, generator=g)
    static = 0.3 + 0.7 * torch.rand(H, W, device=dev, generator=g)
    x = np.arange(11)
    w1 = np.exp(-(x - 5) ** 2 / (2 * 1.5 ** 2)).astype(np.float32)
    w1 = torch.tensor(w1 / w1.sum(), device=dev)
    window = (w1[:, None] @ w1[None, :])[None, None].expand(3, 1, 11, 11).contiguous()

    def torch_step():
        a, b, s = img.clone().requires_grad_(True), gt, static.clone().requires_grad_(True)
        im, gi = a * s, b * s
        conv = lambda t: F.conv2d(t, window, padding=5, groups=3)
        mu1, mu2 = conv(im), conv(gi)
        s1, s2, s12 = conv(im * im) - mu1.pow(2), conv(gi * gi) - mu2.pow(2), conv(im * gi) - mu1 * mu2
        C1, C2 = 0.01 ** 2, 0.03 ** 2
        ssim_map = ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1.pow(2) + mu2.pow(2) + C1) * (s1 + s2 + C2))
        mse = ((im - gi) ** 2).view(3, -1).mean(1, keepdim=True)
        _psnr = (20 * torch.log10(1.0 / torch.sqrt(mse))).mean()
        (((1.0 - 0.2) * torch.abs(im - gi) + 0.2 * (1.0 - ssim_map)).mean()).backward()

    def fused_step():
        a, s = img.clone().requires_grad_(True), static.clone().requires_grad_(True)
        loss, _psnr = photometric_loss(a, gt, s, 0.2)
        loss.backward()

    def timeit(fn):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters

    t_f, t_t = timeit(fused_step), timeit(torch_step)
    print(json.dumps({"H": H, "W": W, "fused_fwd_bwd_ms": round(t_f, 4), "torch_fwd_bwd_ms": round(t_t, 4),
                      "speedup": round(t_t / t_f, 2), "note": "both include two clone() kernels and autograd dispatch"}))


if __name__ == "__main__":
    main()
