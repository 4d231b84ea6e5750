"""Shared helpers of the parity tests: run the same inputs through (a) the new CUDA path via the C ABI / torch binding,
(b) the reference CUDA extension built into oracle/_ref, (c) the C oracle on the CPU, and expose outputs AND the
intermediate arrays (radii, tiles touched, sorted keys, point list, tile ranges, final_T, n_contrib) for comparison.

Tolerances (BASELINE.json north_star): integers bit-exact; forward colour/depth/alpha <= 1e-5 max-abs; gradients
<= 1e-3 relative (atomic order is nondeterministic in the reference).
"""
from __future__ import annotations

import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

FWD_TOL = 1e-5
GRAD_RTOL = 1e-3

LIB_PATH = os.path.join(ROOT, "das3r_b200", "libdas3r_raster.so")


class _Layout(ctypes.Structure):
    _fields_ = [(n, ctypes.c_size_t) for n in (
        "g_recA", "g_recB", "g_recC", "g_tiles_touched", "g_clamped", "geom_bytes",
        "i_final_T", "i_n_contrib", "i_ranges", "i_tile_counts", "i_cursors", "i_totals", "image_bytes",
        "b_unsorted", "b_point_list", "b_keys_sorted", "b_rect_off", "b_rect_cnt", "b_rect_list", "b_rect_gid",
        "binning_bytes")]


_cabi = None


def cabi():
    """The C-ABI shared library loaded with ctypes (no torch in between)."""
    global _cabi
    if _cabi is None:
        lib = ctypes.CDLL(LIB_PATH)
        lib.das3r_last_error.restype = ctypes.c_char_p
        lib.das3r_raster_backward_scratch_bytes.restype = ctypes.c_size_t
        lib.das3r_knn_scratch_bytes.restype = ctypes.c_size_t
        _cabi = lib
    return _cabi


def layout(P, W, H, R) -> _Layout:
    L = _Layout()
    rc = cabi().das3r_raster_state_layout_query(ctypes.c_int32(P), ctypes.c_int32(W), ctypes.c_int32(H),
                                                ctypes.c_int32(R), ctypes.byref(L))
    assert rc == 0, cabi().das3r_last_error()
    return L


def to_dev(vi: dict, device) -> dict:
    return {k: (v.to(device) if isinstance(v, torch.Tensor) else v) for k, v in vi.items()}


def _empty():
    return torch.Tensor([])


def _opt(vi, k):
    v = vi.get(k)
    return _empty() if v is None else v


def _settings(mod, vi, debug=False):
    return mod.GaussianRasterizationSettings(
        image_height=int(vi["image_height"]), image_width=int(vi["image_width"]), tanfovx=vi["tanfovx"],
        tanfovy=vi["tanfovy"], bg=vi["bg"], scale_modifier=vi.get("scale_modifier", 1.0), viewmatrix=vi["viewmatrix"],
        projmatrix=vi["projmatrix"], sh_degree=int(vi["sh_degree"]), campos=vi["campos"], prefiltered=False, debug=debug)


def _view(buf: torch.Tensor, off: int, count: int, dtype) -> torch.Tensor:
    nbytes = count * torch.empty((), dtype=dtype).element_size()
    return buf[off:off + nbytes].view(dtype)


def run_ours(vi: dict, dL=None, debug=True, extras=True) -> dict:
    """New CUDA path through the drop-in Python surface (autograd) + intermediates from the state blobs."""
    import das3r_b200.rasterizer as R

    P = vi["means3D"].shape[0]
    W, H = int(vi["image_width"]), int(vi["image_height"])
    leaf = {}
    for k in ("means3D", "opacities", "shs", "colors_precomp", "scales", "rotations", "cov3D_precomp"):
        v = vi.get(k)
        leaf[k] = None if v is None else v.detach().clone().requires_grad_(True)
    means2D = torch.zeros(P, 3, device=vi["means3D"].device, requires_grad=True)
    rast = R.GaussianRasterizer(_settings(R, vi, debug))
    res = rast(means3D=leaf["means3D"], means2D=means2D, opacities=leaf["opacities"], shs=leaf["shs"],
               colors_precomp=leaf["colors_precomp"], scales=leaf["scales"], rotations=leaf["rotations"],
               cov3D_precomp=leaf["cov3D_precomp"], return_depth_alpha=extras)
    color, radii = res[0], res[1]
    out = dict(color=color.detach(), radii=radii.detach())
    if extras:
        out["depth"], out["alpha"] = res[2].detach(), res[3].detach()
    fn = color.grad_fn
    if fn is not None and P > 0:
        saved = fn.saved_tensors
        geom, binning, img = saved[7], saved[8], saved[9]
        Rn = fn.num_rendered
        L = layout(P, W, H, Rn)
        T = ((W + 15) // 16) * ((H + 15) // 16)
        rec = _view(geom, L.g_recA, 12 * P, torch.float32).view(P, 12)
        out.update(num_rendered=Rn, rec=rec, tiles_touched=_view(geom, L.g_tiles_touched, P, torch.int32),
                   clamped=_view(geom, L.g_clamped, P, torch.uint8),
                   final_T=_view(img, L.i_final_T, W * H, torch.float32),
                   n_contrib=_view(img, L.i_n_contrib, W * H, torch.int32),
                   ranges=_view(img, L.i_ranges, 2 * T, torch.int32).view(T, 2))
        if Rn > 0:
            out["point_list"] = _view(binning, L.b_point_list, Rn, torch.int32)
            if debug:
                out["keys_sorted"] = _view(binning, L.b_keys_sorted, Rn, torch.int64)
    if dL is not None and fn is not None:
        color.backward(dL)
        out["grads"] = dict(
            dL_dmeans3D=leaf["means3D"].grad, dL_dmeans2D=means2D.grad, dL_dopacity=leaf["opacities"].grad,
            dL_dsh=None if leaf["shs"] is None else leaf["shs"].grad,
            dL_dcolors=None if leaf["colors_precomp"] is None else leaf["colors_precomp"].grad,
            dL_dscales=None if leaf["scales"] is None else leaf["scales"].grad,
            dL_drotations=None if leaf["rotations"] is None else leaf["rotations"].grad,
            dL_dcov3D=None if leaf["cov3D_precomp"] is None else leaf["cov3D_precomp"].grad)
    return out


def _obtain(off, count, itemsize, align=128):
    off = (off + align - 1) & ~(align - 1)
    return off, off + count * itemsize


def run_ref(vi: dict, dL=None, debug=False) -> dict:
    """Reference CUDA rasterizer (oracle/_ref) through its own pybind API + intermediates replayed from its byte
    buffers (GeometryState/BinningState/ImageState::fromChunk, rasterizer_impl.cu:155-194)."""
    import oracle

    C, _ = oracle.load_reference()
    P = vi["means3D"].shape[0]
    W, H = int(vi["image_width"]), int(vi["image_height"])
    args = (vi["bg"], vi["means3D"], _opt(vi, "colors_precomp"), vi["opacities"], _opt(vi, "scales"),
            _opt(vi, "rotations"), vi.get("scale_modifier", 1.0), _opt(vi, "cov3D_precomp"), vi["viewmatrix"],
            vi["projmatrix"], vi["tanfovx"], vi["tanfovy"], H, W, _opt(vi, "shs"), int(vi["sh_degree"]), vi["campos"],
            False, debug)
    Rn, color, radii, geom, binning, img = C.rasterize_gaussians(*args)
    out = dict(color=color, radii=radii, num_rendered=Rn)
    if P > 0:
        o = 0
        o, e = _obtain(o, P, 4); out["depths"] = _view(geom, o, P, torch.float32); o = e
        o, e = _obtain(o, 3 * P, 1); out["clamped"] = _view(geom, o, 3 * P, torch.uint8).view(P, 3); o = e
        o, e = _obtain(o, P, 4); o = e  # internal_radii
        o, e = _obtain(o, P, 8); out["xy"] = _view(geom, o, 2 * P, torch.float32).view(P, 2); o = e
        o, e = _obtain(o, 6 * P, 4); out["cov3D"] = _view(geom, o, 6 * P, torch.float32).view(P, 6); o = e
        o, e = _obtain(o, P, 16); out["conic_opacity"] = _view(geom, o, 4 * P, torch.float32).view(P, 4); o = e
        o, e = _obtain(o, 3 * P, 4); out["rgb"] = _view(geom, o, 3 * P, torch.float32).view(P, 3); o = e
        o, e = _obtain(o, P, 4); out["tiles_touched"] = _view(geom, o, P, torch.int32); o = e
        N = W * H
        T = ((W + 15) // 16) * ((H + 15) // 16)
        o = 0
        o, e = _obtain(o, N, 4); out["final_T"] = _view(img, o, N, torch.float32); o = e
        o, e = _obtain(o, N, 4); out["n_contrib"] = _view(img, o, N, torch.int32); o = e
        o, e = _obtain(o, N, 8); out["ranges"] = _view(img, o, 2 * T, torch.int32).view(T, 2); o = e
        if Rn > 0:
            o = 0
            o, e = _obtain(o, Rn, 4); out["point_list"] = _view(binning, o, Rn, torch.int32); o = e
            o, e = _obtain(o, Rn, 4); o = e
            o, e = _obtain(o, Rn, 8); out["keys_sorted"] = _view(binning, o, Rn, torch.int64); o = e
    if dL is not None:
        bargs = (vi["bg"], vi["means3D"], radii, _opt(vi, "colors_precomp"), _opt(vi, "scales"), _opt(vi, "rotations"),
                 vi.get("scale_modifier", 1.0), _opt(vi, "cov3D_precomp"), vi["viewmatrix"], vi["projmatrix"],
                 vi["tanfovx"], vi["tanfovy"], dL, _opt(vi, "shs"), int(vi["sh_degree"]), vi["campos"], geom, Rn,
                 binning, img, debug)
        g = C.rasterize_gaussians_backward(*bargs)
        names = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
                 "dL_drotations")
        out["grads"] = dict(zip(names, g))
    return out


def run_oracle(vi: dict, dL=None) -> dict:
    """C oracle on the CPU (inputs are moved to the host)."""
    import oracle

    cpu = {k: (v.detach().cpu() if isinstance(v, torch.Tensor) else v) for k, v in vi.items()}
    fw = oracle.forward(**cpu)
    out = dict(fw)
    if dL is not None:
        out["grads"] = oracle.backward(fw, dL.detach().cpu(), **cpu)
    return out


def np_(t):
    if t is None:
        return None
    if isinstance(t, torch.Tensor):
        return t.detach().cpu().numpy()
    return np.asarray(t)


def rel_err(a, b) -> float:
    """max |a-b| / max|b| — the gradient metric (relative to the largest reference entry of that tensor)."""
    a, b = np_(a).astype(np.float64), np_(b).astype(np.float64)
    if a.size == 0:
        return 0.0
    denom = max(np.abs(b).max(), 1e-30)
    return float(np.abs(a - b).max() / denom)


def compare_forward(ours: dict, ref: dict, P: int, has_sh: bool = True) -> dict:
    """Mismatch counts / max errors between the new path and a reference-layout dict (run_ref or run_oracle).
    has_sh=False (colors_precomp given): the reference never writes its rgb / clamped arrays, so they are not compared."""
    r = {}
    r["radii_mismatch"] = int((np_(ours["radii"]) != np_(ref["radii"])).sum())
    r["color_maxabs"] = float(np.abs(np_(ours["color"]) - np_(ref["color"])).max())
    if P == 0 or "tiles_touched" not in ours:
        return r
    r["num_rendered"] = (int(ours["num_rendered"]), int(ref["num_rendered"]))
    r["tiles_mismatch"] = int((np_(ours["tiles_touched"]).astype(np.int64) != np_(ref["tiles_touched"]).astype(np.int64)).sum())
    vis = np_(ref["radii"]) > 0
    rec = np_(ours["rec"])
    r["xy_mismatch"] = int((rec[vis, 0:2] != np_(ref["xy"])[vis]).sum())
    r["depth_mismatch"] = int((rec[vis, 11] != np_(ref["depths"])[vis]).sum())
    r["conic_mismatch"] = int((rec[vis, 4:8] != np_(ref["conic_opacity"])[vis]).sum())
    r["rgb_maxabs"] = 0.0
    if has_sh:
        rgb_ref = np_(ref["rgb"])[vis]
        r["rgb_maxabs"] = float(np.abs(rec[vis, 8:11] - rgb_ref).max()) if vis.any() else 0.0
        cl = np_(ours["clamped"])[vis]
        clr = np_(ref["clamped"])[vis]
        ours_bits = np.stack([(cl >> c) & 1 for c in range(3)], axis=1)
        r["clamped_mismatch"] = int((ours_bits != (clr != 0)).sum())
    r["ranges_mismatch"] = int((np_(ours["ranges"]).astype(np.int64) != np_(ref["ranges"]).astype(np.int64)).sum())
    if r["num_rendered"][0] == r["num_rendered"][1] and r["num_rendered"][0] > 0:
        r["point_list_mismatch"] = int((np_(ours["point_list"]).astype(np.int64) !=
                                        np_(ref["point_list"]).astype(np.int64)).sum())
        if "keys_sorted" in ours:
            r["keys_mismatch"] = int((np_(ours["keys_sorted"]).astype(np.int64) !=
                                      np_(ref["keys_sorted"]).astype(np.int64)).sum())
    r["final_T_maxabs"] = float(np.abs(np_(ours["final_T"]) - np_(ref["final_T"])).max())
    r["n_contrib_mismatch"] = int((np_(ours["n_contrib"]).astype(np.int64) != np_(ref["n_contrib"]).astype(np.int64)).sum())
    return r


def rel_err_elementwise(a, b, floor=0.1) -> float:
    """max over the entries with |b| >= floor * max|b| of |a-b| / |b|: a true per-element relative error, restricted to
    entries that are not the residue of a cancellation (for those only the tensor-wide metric above is meaningful)."""
    a, b = np_(a).astype(np.float64), np_(b).astype(np.float64)
    if a.size == 0:
        return 0.0
    big = np.abs(b) >= floor * max(np.abs(b).max(), 1e-30)
    if not big.any():
        return 0.0
    return float((np.abs(a - b)[big] / np.abs(b)[big]).max())


def compare_grads(ours: dict, ref: dict) -> dict:
    """Per gradient tensor: max|a-b| / max|b| (key = the tensor's name) and the element-wise relative error of its
    significant entries (key = name + ':elementwise'). Both are held to GRAD_RTOL by the callers."""
    r = {}
    for k, v in ours.items():
        if v is None or k not in ref or ref[k] is None:
            continue
        a, b = np_(v), np_(ref[k])
        if a.size == 0:
            continue
        r[k] = rel_err(a.reshape(b.shape), b)
        r[k + ":elementwise"] = rel_err_elementwise(a.reshape(b.shape), b)
    return r
