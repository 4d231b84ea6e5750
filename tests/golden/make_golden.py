"""Generates the golden vectors of tests/golden/*.npz by running the REFERENCE CUDA rasterizer (oracle/_ref, built
from /root/reference by oracle/build_ref.sh: the vendored diff-gaussian-rasterization @59f5f77e with DAS3R's 0.001
near plane, sm_100, CUDA 12.9) on the seeded cases of tests/cases.py. Needs a GPU:

    gpurun -- 'python tests/golden/make_golden.py'      # writes gpurun_out/golden/*.npz
    cp gpurun_out/golden/*.npz tests/golden/

Each file holds the inputs (so the fixture is self-contained) and the reference's outputs and intermediates:
integers (radii, tiles_touched, point_list, sorted keys, ranges, n_contrib), the image, final_T and the 8 gradients.
simple-knn's distCUDA2 on the case's points is stored too.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import _harness as Hn  # noqa: E402
from cases import CASE_NAMES, make_case  # noqa: E402
import oracle  # noqa: E402


def main():
    out_dir = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out_dir, exist_ok=True)
    dev = torch.device("cuda:0")
    _, knn = oracle.load_reference()
    for name in CASE_NAMES:
        vi, dL = make_case(name)
        ref = Hn.run_ref(Hn.to_dev(vi, dev), dL.to(dev))
        torch.cuda.synchronize()
        d = {}
        for k, v in vi.items():
            if isinstance(v, torch.Tensor):
                d["in_" + k] = v.numpy()
            elif v is not None:
                d["in_" + k] = np.asarray(v)
        d["in_dL"] = dL.numpy()
        for k in ("color", "radii", "xy", "depths", "conic_opacity", "rgb", "clamped", "tiles_touched", "final_T",
                  "n_contrib", "ranges", "point_list", "keys_sorted", "cov3D"):
            if k in ref:
                d["out_" + k] = Hn.np_(ref[k])
        d["out_num_rendered"] = np.asarray(ref["num_rendered"])
        for k, v in ref["grads"].items():
            d["grad_" + k] = Hn.np_(v)
        d["out_knn"] = knn.distCUDA2(vi["means3D"].to(dev)).cpu().numpy()
        path = os.path.join(out_dir, name + ".npz")
        np.savez_compressed(path, **d)
        print(name, os.path.getsize(path), "bytes, R =", int(ref["num_rendered"]))


if __name__ == "__main__":
    main()
