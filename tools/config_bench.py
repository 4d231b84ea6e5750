"""Developer tool (GPU box): views/s of this implementation and of the reference CUDA rasterizer (oracle/_ref) on the
BASELINE.json configs other than the headline one, single stream, per-view API. C3/C4: forward+backward; C5:
forward only (render.py). Usage: python tools/config_bench.py C3 [iters]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from das3r_b200 import load_native  # noqa: E402
from das3r_b200.synth import CONFIGS, synth_scene, view_inputs, grad_seed  # noqa: E402
import oracle  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C3"
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 12
    fwd_only = name == "C5"
    dev = torch.device("cuda:0")
    kw = dict(CONFIGS[name])
    sc = synth_scene(**kw)
    nposes = kw.get("n_poses") or sc.V
    views = [int(i * (nposes - 1) / 3) for i in range(4)]
    pool = []
    for v in views:
        vi = view_inputs(sc, v)
        pool.append({k: (x.contiguous().to(dev) if isinstance(x, torch.Tensor) else x) for k, x in vi.items()})
    dL = grad_seed(sc).to(dev)
    e = torch.Tensor([])
    res = {"config": name, "P": kw["P"], "W": kw["W"], "H": kw["H"], "mode": "forward only" if fwd_only else "forward+backward"}

    def bench(C):
        def run(vi):
            R, color, radii, geom, binning, img = C.rasterize_gaussians(
                vi["bg"], vi["means3D"], e, vi["opacities"], vi["scales"], vi["rotations"], 1.0, e, vi["viewmatrix"],
                vi["projmatrix"], vi["tanfovx"], vi["tanfovy"], vi["image_height"], vi["image_width"], vi["shs"], 3,
                vi["campos"], False, False)
            if not fwd_only:
                C.rasterize_gaussians_backward(
                    vi["bg"], vi["means3D"], radii, e, vi["scales"], vi["rotations"], 1.0, e, vi["viewmatrix"],
                    vi["projmatrix"], vi["tanfovx"], vi["tanfovy"], dL, vi["shs"], 3, vi["campos"], geom, R, binning,
                    img, False)
            return R

        for i in range(3):
            R = run(pool[i % len(pool)])
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(iters):
            run(pool[i % len(pool)])
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / iters, R

    ms, R = bench(load_native())
    res.update(ours_ms_per_view=round(ms, 4), ours_views_per_s=round(1e3 / ms, 1), num_rendered=int(R))
    if oracle.reference_available():
        ref, _ = oracle.load_reference()
        ms_r, _ = bench(ref)
        res.update(ref_ms_per_view=round(ms_r, 4), ref_views_per_s=round(1e3 / ms_r, 1), speedup=round(ms_r / ms, 2))
    print(json.dumps(res))


if __name__ == "__main__":
    main()
