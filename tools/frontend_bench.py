"""Developer tool (GPU box): the fused DAS3R render front-end (das3r_frontend_forward / _backward) against the
reference's torch formulation of the same lines (gaussian_renderer/__init__.py:83-126 restated with torch ops) at
P Gaussians and V views. Prints one JSON line. Usage: python tools/frontend_bench.py [P] [V] [iters]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from das3r_b200 import load_native  # noqa: E402


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    V = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    iters = int(sys.argv[3]) if len(sys.argv) > 3 else 20
    C = load_native()
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    rn = lambda *s: torch.randn(*s, device=dev, generator=g)
    n_pix = 2 * P
    m = dict(xyz=rn(P, 3), rotation=rn(P, 4), opacity=rn(P, 1), scaling=-3 + 0.5 * rn(P, 3), features_dc=rn(P, 1, 3),
             features_rest=0.1 * rn(P, 15, 3), conf_static=torch.rand(n_pix, device=dev, generator=g))
    ci = torch.sort(torch.randperm(n_pix, device=dev, generator=g)[:P]).values
    poses = torch.cat([torch.tensor([1.0, 0, 0, 0], device=dev) + 0.2 * rn(V, 4), 0.3 * rn(V, 3)], dim=1)
    gin = dict(dm=[rn(P, 3) for _ in range(V)], dr=[rn(P, 4) for _ in range(V)], do=rn(P, 1), ds=rn(P, 3), dsh=rn(P, 16, 3))
    pad = torch.empty(512 << 20, dtype=torch.uint8, device=dev)  # L2 flush between iterations

    def fused_fwd():
        return C.frontend_forward(m["xyz"], m["rotation"], m["opacity"], m["scaling"], m["features_dc"], m["features_rest"],
                                  m["conf_static"], ci, poses)

    def fused_bwd():
        return C.frontend_backward(m["xyz"], m["rotation"], m["opacity"], m["scaling"], m["conf_static"], ci, poses,
                                   gin["do"], gin["ds"], gin["dsh"], gin["dm"], gin["dr"])

    def torch_fwd(leaves, pl):
        means, rots = [], []
        for v in range(V):
            q = pl[v, :4]
            r, x, y, z = q / torch.sqrt((q * q).sum())
            w2c = torch.eye(4, device=dev)
            R = torch.stack([torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)]),
                             torch.stack([2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)]),
                             torch.stack([2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)])])
            w2c = torch.cat([torch.cat([R, pl[v, 4:].reshape(3, 1)], dim=1), w2c[3:]], dim=0)
            xyz = leaves["xyz"].clone()
            rot = leaves["rotation"].clone()
            homo = torch.cat((xyz, torch.ones(P, 1, device=dev)), dim=1)
            means.append((w2c @ homo.T).T[:, :3])
            w1, x1, y1, z1 = q.unbind(-1)
            w2, x2, y2, z2 = rot.unbind(-1)
            rots.append(torch.stack([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                                     w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2], dim=-1))
        opac = torch.sigmoid(leaves["opacity"]) * leaves["conf_static"].reshape(-1, 1)[ci]
        return means, rots, opac, torch.exp(leaves["scaling"]), torch.cat((leaves["features_dc"], leaves["features_rest"]), dim=1)

    def torch_step():
        leaves = {k: v.detach().requires_grad_(True) for k, v in m.items()}
        pl = poses.detach().requires_grad_(True)
        means, rots, opac, scales, shs = torch_fwd(leaves, pl)
        torch.autograd.backward(means + rots + [opac, scales, shs], gin["dm"] + gin["dr"] + [gin["do"], gin["ds"], gin["dsh"]])

    def timeit(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        tot = 0.0
        for _ in range(iters):
            pad.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            tot += a.elapsed_time(b)
        return tot / iters

    t_f, t_b, t_ref = timeit(fused_fwd), timeit(fused_bwd), timeit(torch_step)
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0
    bytes_f = P * (248 + 208 + 28 * V)   # reads (incl. the 8-byte conf index) + writes, see DESIGN.md section 7
    bytes_b = P * (264 + 240 + 28 * V)
    print(json.dumps({"P": P, "V": V, "fused_fwd_ms": round(t_f, 4), "fused_bwd_ms": round(t_b, 4),
                      "torch_fwd_bwd_ms": round(t_ref, 4), "speedup": round(t_ref / (t_f + t_b), 2),
                      "fwd_gbs": round(bytes_f / t_f / 1e6, 1), "bwd_gbs": round(bytes_b / t_b / 1e6, 1),
                      "fwd_frac_of_hbm_peak": round(bytes_f / t_f / 1e6 / peak, 3),
                      "bwd_frac_of_hbm_peak": round(bytes_b / t_b / 1e6 / peak, 3), "l2": "flushed between iterations"}))


if __name__ == "__main__":
    main()
