"""Developer tool (GPU box): render.py's all-frames loop (render.py:74-124: for every pose render one frame, forward
only, and hand the image to the host) on the C5-shaped synthetic scene (2 M Gaussians, 854x480, poses interpolated along
the camera arc). Two arms on the same frames:

  ours       MultiViewRasterizer under torch.no_grad(): batches of frames, their forwards spread over CUDA streams,
             every image copied to pinned host memory
  reference  the reference CUDA rasterizer (oracle/_ref), one frame at a time as render.py does, same copy

With --gpus N semantics left to sharding.shard_frames (no communication), this tool times one GPU's share.
Usage: python tools/render_loop_bench.py [frames] [batch]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from das3r_b200 import load_native  # noqa: E402
from das3r_b200.multiview import MultiViewRasterizer  # noqa: E402
from das3r_b200.rasterizer import GaussianRasterizationSettings  # noqa: E402
from das3r_b200.synth import CONFIGS, synth_scene, view_inputs  # noqa: E402
import oracle  # noqa: E402


def main():
    frames = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    dev = torch.device("cuda:0")
    kw = dict(CONFIGS["C5"])
    sc = synth_scene(**kw)
    W, H, nposes = kw["W"], kw["H"], kw["n_poses"]
    pool = []
    for i in range(8):  # 8 distinct poses along the path, cycled (each is a full set of posed means / rotations)
        vi = view_inputs(sc, int(i * (nposes - 1) / 7))
        pool.append({k: (x.contiguous().to(dev) if isinstance(x, torch.Tensor) else x) for k, x in vi.items()})
    host = [torch.empty(3, H, W).pin_memory() for _ in range(batch)]
    e = torch.Tensor([])

    def settings(vi):
        return GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=vi["tanfovx"], tanfovy=vi["tanfovy"], bg=vi["bg"], scale_modifier=1.0,
            viewmatrix=vi["viewmatrix"], projmatrix=vi["projmatrix"], sh_degree=3, campos=vi["campos"], prefiltered=False,
            debug=False)

    mv = MultiViewRasterizer(streams=3)

    def ours():
        with torch.no_grad():
            for f0 in range(0, frames, batch):
                vs = [pool[(f0 + j) % len(pool)] for j in range(min(batch, frames - f0))]
                colors, _ = mv([settings(v) for v in vs], [v["means3D"] for v in vs], [v["rotations"] for v in vs],
                               vs[0]["opacities"], vs[0]["shs"], vs[0]["scales"])
                for j, c in enumerate(colors):
                    host[j].copy_(c, non_blocking=True)
                torch.cuda.current_stream().synchronize()  # the frames are on the host (render.py writes them out)

    def reference(C):
        def run():
            for f in range(frames):
                vi = pool[f % len(pool)]
                out = C.rasterize_gaussians(vi["bg"], vi["means3D"], e, vi["opacities"], vi["scales"], vi["rotations"], 1.0,
                                            e, vi["viewmatrix"], vi["projmatrix"], vi["tanfovx"], vi["tanfovy"], H, W,
                                            vi["shs"], 3, vi["campos"], False, False)
                host[0].copy_(out[1], non_blocking=True)
                torch.cuda.current_stream().synchronize()
        return run

    def timeit(fn):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)

    ms = timeit(ours)
    res = {"config": "C5", "P": kw["P"], "W": W, "H": H, "frames": frames, "batch": batch,
           "ours_frames_per_s": round(frames / ms * 1e3, 1), "d2h_bytes_per_frame": 3 * H * W * 4}
    ms1 = timeit(reference(load_native()))
    res["ours_per_frame_api_frames_per_s"] = round(frames / ms1 * 1e3, 1)
    if oracle.reference_available():
        ms_r = timeit(reference(oracle.load_reference()[0]))
        res.update(ref_frames_per_s=round(frames / ms_r * 1e3, 1), speedup=round(ms_r / ms, 2))
    print(json.dumps(res))


if __name__ == "__main__":
    main()
