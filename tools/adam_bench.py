"""Developer tool (GPU box): one optimizer step over DAS3R's seven parameter groups at P Gaussians - FusedAdam against
torch.optim.Adam (default multi-tensor path, as the reference constructs it) and torch's own fused=True variant.
Usage: python tools/adam_bench.py [P] [iters]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from das3r_b200.optim import FusedAdam  # noqa: E402


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    dev = torch.device("cuda:0")
    shapes = [(P, 3), (P, 1, 3), (P, 15, 3), (P, 1), (P, 3), (P, 4), (50, 288, 512)]
    lrs = [1.6e-4, 2.5e-3, 2.5e-3 / 20, 0.05, 5e-3, 1e-3, 3e-3]
    n_el = sum(int(torch.Size(s).numel()) for s in shapes)
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0

    def run(make):
        ps = [torch.randn(*s, device=dev).requires_grad_(True) for s in shapes]
        for p in ps:
            p.grad = torch.randn_like(p)
        opt = make([{"params": [p], "lr": lr} for p, lr in zip(ps, lrs)])
        for _ in range(3):
            opt.step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            opt.step()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters

    t_ours = run(lambda l: FusedAdam(l, lr=0.0, eps=1e-15))
    t_ref = run(lambda l: torch.optim.Adam(l, lr=0.0, eps=1e-15))
    t_lib = run(lambda l: torch.optim.Adam(l, lr=0.0, eps=1e-15, fused=True))
    gbs = n_el * 28 / t_ours / 1e6
    print(json.dumps({"P": P, "elements": n_el, "fused_ms": round(t_ours, 4), "torch_adam_ms": round(t_ref, 4),
                      "torch_adam_fused_ms": round(t_lib, 4), "speedup_vs_reference": round(t_ref / t_ours, 2),
                      "gbs": round(gbs, 1), "frac_of_hbm_peak": round(gbs / peak, 3),
                      "note": "parameters + moments + gradients = 1.1 GB > L2, every step streams them from HBM"}))


if __name__ == "__main__":
    main()
