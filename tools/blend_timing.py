"""Developer tool (GPU box, with a -DDAS3R_BLEND_TIMING variant library installed): per-warp start/end/cycles of
blend_fwd for one C2 view."""
import ctypes, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from das3r_b200 import load_native
from das3r_b200.synth import CONFIGS, synth_scene, view_inputs
C = load_native()
lib = ctypes.CDLL(os.path.join(ROOT, "das3r_b200", "libdas3r_raster.so"))
dev = torch.device("cuda:0")
kw = dict(CONFIGS["C2"]); kw.pop("n_poses", None)
sc = synth_scene(**kw)
vi = {k: (x.contiguous().to(dev) if isinstance(x, torch.Tensor) else x) for k, x in view_inputs(sc, 0, sh_degree=3).items()}
e = torch.Tensor([])
for _ in range(3):
    out = C.rasterize_gaussians(vi["bg"], vi["means3D"], e, vi["opacities"], vi["scales"], vi["rotations"], 1.0, e, vi["viewmatrix"],
        vi["projmatrix"], vi["tanfovx"], vi["tanfovy"], vi["image_height"], vi["image_width"], vi["shs"], 3, vi["campos"], False, False)
torch.cuda.synchronize()
nw = 4608
buf = np.zeros(4 * nw, np.uint64)
rc = lib.das3r_debug_blend_timing(buf.ctypes.data_as(ctypes.c_void_p), 4 * nw)
assert rc == 0
b = buf.reshape(nw, 4)
t0 = b[:, 0].astype(np.int64); t1 = b[:, 1].astype(np.int64); cyc = b[:, 2].astype(np.int64)
cnt = (b[:, 3] & np.uint64(0xffffff)).astype(np.int64); smid = (b[:, 3] >> np.uint64(48)).astype(np.int64)
base = t0.min()
print("kernel span ns", t1.max() - base, "start spread ns", t0.max() - base)
dur = t1 - t0
order = np.argsort(-dur)
print("dur ns: mean", dur.mean(), "p50", np.percentile(dur, 50), "p90", np.percentile(dur, 90), "p99", np.percentile(dur, 99), "max", dur.max())
print("cnt: mean", cnt.mean(), "max", cnt.max())
print("slowest warps: (dur ns, cycles, cnt, cycles/entry, start, sm)")
for i in order[:12]:
    print(dur[i], cyc[i], cnt[i], round(cyc[i] / max(cnt[i], 1), 1), t0[i] - base, smid[i])
# linear fit cycles ~ a + b*cnt
A = np.vstack([np.ones_like(cnt), cnt]).T.astype(np.float64)
coef, *_ = np.linalg.lstsq(A, cyc.astype(np.float64), rcond=None)
print("fit cycles = %.0f + %.1f * cnt" % (coef[0], coef[1]))
# per-SM finish times
fin = np.array([t1[smid == s].max() - base for s in np.unique(smid)])
print("per-SM finish ns: min", fin.min(), "mean", fin.mean(), "max", fin.max())

wpall = np.zeros(8 * 65536, np.uint64)
if hasattr(lib, "das3r_debug_walk_prof") and lib.das3r_debug_walk_prof(wpall.ctypes.data_as(ctypes.c_void_p), 8 * 65536) == 0:
    wp = wpall[:8 * nw].reshape(nw, 8).astype(np.int64)
    wq = wpall[8 * 32768:8 * 32768 + 8 * nw].reshape(nw, 8).astype(np.int64)
    names = ["bookkeeping", "start", "land", "blend burst"]
    for label, sel in (("all warps", slice(None)), ("12 slowest", order[:12])):
        tot = wp[sel, :4].sum(0)
        rounds, lanes, lands, bursts = wp[sel, 4].sum(), wp[sel, 5].sum(), wp[sel, 6].sum(), wp[sel, 7].sum()
        ent = cnt[sel].sum()
        print(label, "cycle share by phase:", dict(zip(names, np.round(tot / tot.sum(), 3))))
        print(label, "rounds per chunk %.2f, active lanes per round %.1f, cycles per round %.0f, bursts per chunk %.2f, lands per chunk %.2f"
              % (rounds / (ent / 32), lanes / max(rounds, 1), tot[3] / max(rounds, 1), bursts / (ent / 32), lands / (ent / 32)))
        q = wq[sel, :6].sum(0)
        print(label, "cycles per round inside the burst: pop+addr %.0f, prep %.0f, apply0 %.0f, apply1 %.0f, keep %.0f, vote %.0f"
              % tuple(q / max(rounds, 1)))
