"""Developer tool (GPU box): DAS3R's training iteration through das3r_b200.train.Trainer on the C2-shaped synthetic
scene (1 M Gaussians, 50 views of 512x288, one random view per iteration, SH degree 0 as in iterations 1-2999), in two
arms that start from the same parameters and see the same views: this repository's fused components, and the reference's
formulation (torch glue + the reference CUDA rasterizer + torch loss + torch.optim.Adam; hooks of tests/test_gpu_train.py).
Prints one JSON line. Usage: python tools/trainer_bench.py [iterations] [sh_degree]"""
import json
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from das3r_b200.synth import CONFIGS, projection_matrix, synth_scene  # noqa: E402
from das3r_b200.train import Camera, GaussianState, OptimizationParams, Trainer, ViewGeometry  # noqa: E402
import oracle  # noqa: E402


def main():
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 100
    D = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    dev = torch.device("cuda:0")
    sc = synth_scene(**{**CONFIGS["C2"], "sh_degree": D})
    P, V, W, H = sc.P, sc.V, sc.W, sc.H
    g = torch.Generator().manual_seed(0)
    conf_index = torch.randperm(V * H * W, generator=g)[:P].sort().values.to(dev)

    def state():
        return GaussianState(
            xyz=sc.xyz.to(dev).clone(), features_dc=sc.shs[:, :1].to(dev).contiguous(),
            features_rest=sc.shs[:, 1:].to(dev).contiguous(), opacity=torch.logit(sc.opacity.clamp(1e-4, 1 - 1e-4)).to(dev),
            scaling=torch.log(sc.scales).to(dev), rotation=sc.rot.to(dev).clone(),
            conf_static=(0.5 + 0.5 * torch.rand(V, H, W, generator=torch.Generator().manual_seed(1))).to(dev),
            Q=sc.poses[:, :4].to(dev).contiguous(), T=sc.poses[:, 4:].to(dev).contiguous(), conf_index=conf_index,
            active_sh_degree=D)

    proj = projection_matrix(0.01, 100.0, 2 * math.atan(sc.tanfovx), 2 * math.atan(sc.tanfovy)).transpose(0, 1).to(dev).contiguous()
    geom = ViewGeometry(image_height=H, image_width=W, tanfovx=sc.tanfovx, tanfovy=sc.tanfovy, projmatrix=proj,
                        bg=torch.zeros(3, device=dev))
    cams = [Camera(uid=v, original_image=torch.rand(3, H, W, generator=g).to(dev)) for v in range(V)]
    opt = OptimizationParams(iterations=4000)

    def run(tr):
        for _ in range(10):
            tr.step()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            tr.step()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / iters, [h[2] for h in tr.history]

    ms_ours, l_ours = run(Trainer(state(), cams, geom, opt, seed=3))
    res = {"P": P, "views": V, "W": W, "H": H, "sh_degree": D, "iterations": iters, "ours_ms_per_iteration": round(ms_ours, 4),
           "ours_iterations_per_s": round(1e3 / ms_ours, 1)}
    if oracle.reference_available():
        from test_gpu_train import _reference_hooks
        rf, lf, adam = _reference_hooks()
        ms_ref, l_ref = run(Trainer(state(), cams, geom, opt, seed=3, render_fn=rf, loss_fn=lf, adam_cls=adam))
        rel = max(abs(a - b) / b for a, b in zip(l_ours, l_ref))
        res.update(ref_ms_per_iteration=round(ms_ref, 4), ref_iterations_per_s=round(1e3 / ms_ref, 1),
                   speedup=round(ms_ref / ms_ours, 2), max_relative_loss_difference=rel, first_losses_ours=l_ours[:3],
                   first_losses_ref=l_ref[:3])
    print(json.dumps(res))


if __name__ == "__main__":
    main()
