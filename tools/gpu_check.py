"""Developer diagnostic (GPU box): parity of the new path vs oracle/_ref and the C oracle on a few scenes, verbose,
plus a rough per-call timing of both CUDA implementations. Not part of the test suite."""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _harness as Hn  # noqa: E402
from das3r_b200.synth import synth_scene, view_inputs, grad_seed  # noqa: E402


def timeit(fn, warm=3, it=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(it):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / it


def main():
    dev = torch.device("cuda:0")
    print(torch.cuda.get_device_name(0))
    cases = [("tiny", dict(P=2000, V=4, W=64, H=48, seed=7), 1, 3),
             ("C1", dict(P=100_000, V=4, W=512, H=288, seed=0), 0, 3),
             ("C2", dict(P=1_000_000, V=50, W=512, H=288, seed=1), 25, 3),
             ("C2-D0", dict(P=1_000_000, V=50, W=512, H=288, seed=1), 25, 0)]
    if len(sys.argv) > 1:
        cases = [c for c in cases if c[0] in sys.argv[1:]]
    for name, kw, v, D in cases:
        sc = synth_scene(**kw)
        vi = Hn.to_dev(view_inputs(sc, v, sh_degree=D), dev)
        vi["means3D"] = vi["means3D"].contiguous()
        dL = grad_seed(sc).to(dev)
        P = kw["P"]
        ours = Hn.run_ours(vi, dL, debug=True)
        ref = Hn.run_ref(vi, dL)
        torch.cuda.synchronize()
        print(f"== {name}: P={P} R={ours.get('num_rendered')} (ref {ref['num_rendered']})")
        print("  vs ref fwd :", json.dumps(Hn.compare_forward(ours, ref, P)))
        print("  vs ref grad:", json.dumps(Hn.compare_grads(ours["grads"], ref["grads"])))
        if P <= 100_000:
            orc = Hn.run_oracle(vi, dL)
            print("  vs C-oracle fwd :", json.dumps(Hn.compare_forward(ours, orc, P)))
            print("  vs C-oracle grad:", json.dumps(Hn.compare_grads(ours["grads"], orc["grads"])))
            print("  ref vs C-oracle grad:", json.dumps(Hn.compare_grads(ref["grads"], orc["grads"])))
        if name == "tiny":
            import numpy as np
            rec = Hn.np_(ours["rec"]); rc = Hn.np_(ref["conic_opacity"]); vis = Hn.np_(ref["radii"]) > 0
            bad = np.where(vis & (rec[:, 4:7] != rc[:, 0:3]).any(axis=1))[0][:5]
            for b in bad:
                print("   conic idx", b, "ours", rec[b, 4:7].view(np.uint32), "ref", rc[b, 0:3].view(np.uint32),
                      "oracle", orc["conic_opacity"][b, 0:3].view(np.uint32))
            print("   ref cov3D vs oracle cov3D mismatches:",
                  int((Hn.np_(ref["cov3D"])[vis] != orc["cov3D"][vis]).sum()), "of", int(vis.sum()) * 6)
            bad = np.where(vis & (Hn.np_(ref["cov3D"]) != orc["cov3D"]).any(axis=1))[0][:3]
            for b in bad:
                print("   cov3D idx", b, "ref", Hn.np_(ref["cov3D"])[b].view(np.uint32), "oracle", orc["cov3D"][b].view(np.uint32))
        rng = Hn.np_(ours["ranges"])
        cnt = rng[:, 1] - rng[:, 0]
        print(f"  tile count mean {cnt.mean():.0f} max {cnt.max()}  T mean {float(ours['final_T'].mean()):.3f}")
        # timing: forward+backward through each implementation's pybind functions
        t_ours = timeit(lambda: Hn.run_ours(vi, dL, debug=False, extras=False))
        t_ref = timeit(lambda: Hn.run_ref(vi, dL))
        print(f"  fwd+bwd ms: ours {t_ours:.3f}  ref {t_ref:.3f}  speedup {t_ref / t_ours:.2f}x")
        del ours, ref
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
