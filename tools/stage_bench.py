"""Developer tool (GPU box): per-stage device times of forward+backward on a BASELINE config via the library's event
hooks. Usage: python tools/stage_bench.py [C1|C2|C3] [D] [iters]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from das3r_b200 import load_native  # noqa: E402
from das3r_b200.synth import CONFIGS, synth_scene, view_inputs, grad_seed  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C2"
    D = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    iters = int(sys.argv[3]) if len(sys.argv) > 3 else 20
    C = load_native()
    dev = torch.device("cuda:0")
    kw = dict(CONFIGS[name])
    kw.pop("n_poses", None)
    sc = synth_scene(**kw)
    views = [0, sc.V // 2, sc.V - 1, sc.V // 3]
    pool = []
    for v in views:
        vi = view_inputs(sc, v, sh_degree=D)
        pool.append({k: (x.contiguous().to(dev) if isinstance(x, torch.Tensor) else x) for k, x in vi.items()})
    dL = grad_seed(sc).to(dev)
    e = torch.Tensor([])

    def run(vi):
        R, color, radii, geom, binning, img = C.rasterize_gaussians(
            vi["bg"], vi["means3D"], e, vi["opacities"], vi["scales"], vi["rotations"], 1.0, e, vi["viewmatrix"],
            vi["projmatrix"], vi["tanfovx"], vi["tanfovy"], vi["image_height"], vi["image_width"], vi["shs"], D,
            vi["campos"], False, False)
        C.rasterize_gaussians_backward(
            vi["bg"], vi["means3D"], radii, e, vi["scales"], vi["rotations"], 1.0, e, vi["viewmatrix"],
            vi["projmatrix"], vi["tanfovx"], vi["tanfovy"], dL, vi["shs"], D, vi["campos"], geom, R, binning, img,
            False)
        return R

    for i in range(5):
        R = run(pool[i % len(pool)])
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(iters):
        run(pool[i % len(pool)])
    b.record()
    torch.cuda.synchronize()
    total = a.elapsed_time(b) / iters
    C.prof_collect(True)
    C.prof_enable(True)
    for i in range(iters):
        run(pool[i % len(pool)])
    torch.cuda.synchronize()
    st = {n: round(ms / max(c, 1), 4) for n, ms, c in C.prof_collect(True) if c}
    C.prof_enable(False)
    print(json.dumps({"config": name, "D": D, "R": R, "ms_per_view_wall": round(total, 4),
                      "sum_stage_ms": round(sum(st.values()), 4), "stages": st}))


if __name__ == "__main__":
    main()
