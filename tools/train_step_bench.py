"""Developer tool (GPU box): DAS3R's training iteration (train_gui.py:555-586) on a synthetic C2-shaped scene, one view
per iteration as in the reference loop, in two arms that start from the same parameters:

  ours       render_inputs (fused front-end) -> GaussianRasterizer (this repository) -> photometric_loss (fused) ->
             backward -> FusedAdam.step()
  reference  the reference's formulation: torch glue of gaussian_renderer/__init__.py:83-126 -> the reference CUDA
             rasterizer (oracle/_ref) -> utils/loss_utils.py formulation in torch -> torch.optim.Adam(l, lr=0, eps=1e-15)

Prints one JSON line: iterations/s of both arms and the largest difference of the per-iteration losses over the first
iterations (same views, same order). Usage: python tools/train_step_bench.py [P] [iters]"""
import json
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from das3r_b200 import load_native  # noqa: E402
from das3r_b200.frontend import render_inputs  # noqa: E402
from das3r_b200.loss import photometric_loss  # noqa: E402
from das3r_b200.optim import FusedAdam  # noqa: E402
from das3r_b200.synth import synth_scene, view_inputs  # noqa: E402
import oracle  # noqa: E402

LRS = dict(xyz=1.6e-4, features_dc=2.5e-3, features_rest=2.5e-3 / 20, opacity=0.05, scaling=5e-3, rotation=1e-3,
           conf_static=3e-3)  # arguments/__init__.py:72-83, gaussian_model.py:233
LAMBDA = 0.2


def make_rasterize(C):
    """autograd wrapper with the reference's call signature around a `_C` module (ours or oracle/_ref)."""
    e = torch.Tensor([])

    class Fn(torch.autograd.Function):
        @staticmethod
        def forward(ctx, means3D, shs, opacities, scales, rotations, cam):
            out = C.rasterize_gaussians(cam["bg"], means3D, e, opacities, scales, rotations, 1.0, e, cam["view"],
                                        cam["proj"], cam["tanx"], cam["tany"], cam["H"], cam["W"], shs, 3, cam["campos"],
                                        False, False)
            R, color, radii, geom, binning, img = out[:6]
            ctx.cam, ctx.R = cam, R
            ctx.save_for_backward(means3D, shs, scales, rotations, radii, geom, binning, img)
            return color

        @staticmethod
        def backward(ctx, g):
            means3D, shs, scales, rotations, radii, geom, binning, img = ctx.saved_tensors
            cam = ctx.cam
            (_m2, _dc, d_op, d_m3, _cov, d_sh, d_sc, d_rot) = C.rasterize_gaussians_backward(
                cam["bg"], means3D, radii, e, scales, rotations, 1.0, e, cam["view"], cam["proj"], cam["tanx"], cam["tany"],
                g.contiguous(), shs, 3, cam["campos"], geom, ctx.R, binning, img, False)
            return d_m3, d_sh, d_op, d_sc, d_rot, None

    return Fn.apply


def torch_frontend(m, pose):
    """gaussian_renderer/__init__.py:83-97,:106,:126 with utils/pose_utils.py:36-104"""
    q = pose[:4]
    r, x, y, z = q / torch.sqrt((q * q).sum())
    R = torch.stack([torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)]),
                     torch.stack([2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)]),
                     torch.stack([2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)])])
    w2c = torch.eye(4, device=pose.device)
    w2c = torch.cat([torch.cat([R, pose[4:].reshape(3, 1)], dim=1), w2c[3:]], dim=0)
    xyz, rot = m["xyz"].clone(), m["rotation"].clone()
    homo = torch.cat((xyz, torch.ones(xyz.shape[0], 1, device=xyz.device)), dim=1)
    means = (w2c @ homo.T).T[:, :3]
    w1, x1, y1, z1 = q.unbind(-1)
    w2, x2, y2, z2 = rot.unbind(-1)
    rots = torch.stack([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                        w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2], dim=-1)
    opacity = torch.sigmoid(m["opacity"]) * m["conf_static"].reshape(-1, 1)
    return means, rots, opacity, torch.exp(m["scaling"]), torch.cat((m["features_dc"], m["features_rest"]), dim=1)


def torch_loss(image, gt, static, window):
    """train_gui.py:567-576 with utils/loss_utils.py:36-67"""
    im, g = image * static, gt * static
    conv = lambda t: F.conv2d(t, window, padding=5, groups=3)
    mu1, mu2 = conv(im), conv(g)
    s1, s2, s12 = conv(im * im) - mu1.pow(2), conv(g * g) - mu2.pow(2), conv(im * g) - mu1 * mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim_map = ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1.pow(2) + mu2.pow(2) + C1) * (s1 + s2 + C2))
    mse = ((im - g) ** 2).view(3, -1).mean(1, keepdim=True)
    psnr = (20 * torch.log10(1.0 / torch.sqrt(mse))).mean()
    return ((1.0 - LAMBDA) * torch.abs(im - g) + LAMBDA * (1.0 - ssim_map)).mean(), psnr


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    V, W, H = 50, 512, 288
    dev = torch.device("cuda:0")
    sc = synth_scene(P=P, V=V, W=W, H=H, seed=1, sh_degree=3)
    vi = view_inputs(sc, 0)
    g = torch.Generator().manual_seed(0)
    init = dict(xyz=vi["means3D"].clone(), rotation=vi["rotations"].clone(),
                opacity=torch.logit(vi["opacities"].clamp(1e-4, 1 - 1e-4)), scaling=torch.log(vi["scales"]),
                features_dc=vi["shs"][:, :1].clone(), features_rest=vi["shs"][:, 1:].clone(),
                conf_static=torch.ones(P))
    # per-view data of the loop: small pose perturbations around view 0 (the scene stays in view), ground-truth images,
    # static masks
    n_views = 8
    poses = torch.cat([torch.tensor([1.0, 0, 0, 0]) + 0.002 * torch.randn(n_views, 4, generator=g),
                       0.002 * torch.randn(n_views, 3, generator=g)], dim=1).to(dev)
    gts = [torch.rand(3, H, W, generator=g).to(dev) for _ in range(n_views)]
    statics = [(0.5 + 0.5 * torch.rand(H, W, generator=g)).to(dev) for _ in range(n_views)]
    cam = dict(bg=vi["bg"].to(dev), view=vi["viewmatrix"].to(dev).contiguous(), proj=vi["projmatrix"].to(dev).contiguous(),
               campos=vi["campos"].to(dev), tanx=vi["tanfovx"], tany=vi["tanfovy"], H=H, W=W)
    order = torch.randint(0, n_views, (iters + 8,), generator=g).tolist()
    x = np.arange(11)
    w1 = np.exp(-(x - 5) ** 2 / (2 * 1.5 ** 2)).astype(np.float32)
    w1 = torch.tensor(w1 / w1.sum(), device=dev)
    window = (w1[:, None] @ w1[None, :])[None, None].expand(3, 1, 11, 11).contiguous()

    def arm(kind):
        m = {k: v.to(dev).contiguous().requires_grad_(True) for k, v in init.items()}
        groups = [{"params": [m[k]], "lr": lr, "name": k} for k, lr in LRS.items()]
        if kind == "ours":
            opt = FusedAdam(groups, lr=0.0, eps=1e-15)
            rast = make_rasterize(load_native())
        else:
            opt = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
            rast = make_rasterize(oracle.load_reference()[0])
        losses = []

        def iteration(i):
            v = order[i]
            pose = poses[v]
            if kind == "ours":
                fe = render_inputs(m["xyz"], m["rotation"], m["opacity"], m["scaling"], m["features_dc"],
                                   m["features_rest"], m["conf_static"], None, pose[None])
                image = rast(fe.means3D[0], fe.shs, fe.opacity, fe.scales, fe.rotations[0], cam)
                loss, _psnr = photometric_loss(image, gts[v], statics[v], LAMBDA)
            else:
                means, rots, opac, scales, shs = torch_frontend(m, pose)
                image = rast(means, shs, opac, scales, rots, cam)
                loss, _psnr = torch_loss(image, gts[v], statics[v], window)
            loss.backward()
            opt.step()
            opt.zero_grad(set_to_none=True)
            return loss.detach()

        for i in range(8):
            losses.append(iteration(i))
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(8, 8 + iters):
            iteration(i)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters, [float(l) for l in losses]

    ms_ours, l_ours = arm("ours")
    res = {"P": P, "W": W, "H": H, "iters": iters, "ours_ms_per_iteration": round(ms_ours, 4),
           "ours_iterations_per_s": round(1e3 / ms_ours, 1)}
    if oracle.reference_available():
        ms_ref, l_ref = arm("reference")
        res.update(ref_ms_per_iteration=round(ms_ref, 4), ref_iterations_per_s=round(1e3 / ms_ref, 1),
                   speedup=round(ms_ref / ms_ours, 2),
                   max_loss_difference_first_8_iterations=float(max(abs(a - b) for a, b in zip(l_ours, l_ref))),
                   first_losses_ours=[round(v, 6) for v in l_ours[:4]], first_losses_ref=[round(v, 6) for v in l_ref[:4]])
    print(json.dumps(res))


if __name__ == "__main__":
    main()
