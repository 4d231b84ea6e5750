"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

ctypes/numpy front end of oracle/raster_oracle.c, the CPU restatement of the reference rasterizer, plus loaders for the
reference CUDA extensions built into oracle/_ref. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
reference legs may import this package; das3r_b200 never does (the product has no CPU fallback).

Parity pin: see the header of raster_oracle.c (pinned against oracle/_ref outputs committed under tests/golden/).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_build", "liboracle.so")
REF_DIR = os.path.join(HERE, "_ref")

_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(HERE, "raster_oracle.c")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.run(["make", "-C", HERE, "-B" if force else "-s", "_build/liboracle.so"], check=True,
                       capture_output=True)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(LIB_PATH)
    return _lib


def load_reference():
    """(ref_dgr_C, ref_knn_C): the reference CUDA extensions built by oracle/build_ref.sh, or raise ImportError."""
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    import torch  # noqa: F401  (the extensions link against libtorch)
    import ref_dgr_C
    import ref_knn_C
    return ref_dgr_C, ref_knn_C


def reference_available() -> bool:
    return os.path.exists(os.path.join(REF_DIR, "ref_dgr_C.so")) and os.path.exists(os.path.join(REF_DIR, "ref_knn_C.so"))


def _f(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def _p(a, ct=ctypes.c_float):
    return None if a is None else a.ctypes.data_as(ctypes.POINTER(ct))


def _np(t):
    """torch tensor / array / None -> contiguous float32 numpy array (empty -> None, like the reference's nullptr)."""
    if t is None:
        return None
    if hasattr(t, "detach"):
        t = t.detach().cpu().numpy()
    t = np.asarray(t)
    if t.size == 0:
        return None
    return np.ascontiguousarray(t, dtype=np.float32)


def forward(*, means3D, opacities, viewmatrix, projmatrix, campos, bg, tanfovx, tanfovy, image_height, image_width,
            sh_degree=0, shs=None, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None,
            scale_modifier=1.0, **_ignored) -> dict:
    """Full forward of the restated reference pipeline (Rasterizer::forward, rasterizer_impl.cu:198-336).
    Returns every output and intermediate the parity tests compare."""
    L = lib()
    means = _np(means3D)
    W, H = int(image_width), int(image_height)
    if means is None:  # P == 0: all-zero image, no background (rasterize_points.cu:68,:81)
        gx, gy = (W + 15) // 16, (H + 15) // 16
        return dict(color=np.zeros((3, H, W), np.float32), radii=np.zeros(0, np.int32),
                    depth=np.zeros((H, W), np.float32), alpha=np.zeros((H, W), np.float32), num_rendered=0,
                    xy=np.zeros((0, 2), np.float32), depths=np.zeros(0, np.float32), cov3D=np.zeros((0, 6), np.float32),
                    rgb=np.zeros((0, 3), np.float32), conic_opacity=np.zeros((0, 4), np.float32),
                    tiles_touched=np.zeros(0, np.uint32), clamped=np.zeros((0, 3), np.uint8),
                    point_offsets=np.zeros(0, np.uint32), keys_sorted=np.zeros(0, np.uint64),
                    point_list=np.zeros(0, np.uint32), ranges=np.zeros((gx * gy, 2), np.uint32),
                    final_T=np.ones(W * H, np.float32), n_contrib=np.zeros(W * H, np.uint32), W=W, H=H, M=0)
    P = means.shape[0]
    shs_ = _np(shs)
    M = shs_.shape[1] if shs_ is not None else 0
    colors_ = _np(colors_precomp)
    scales_, rots_, cov_ = _np(scales), _np(rotations), _np(cov3D_precomp)
    opac = _np(opacities).reshape(-1)
    view, proj, cam, bg_ = _np(viewmatrix).reshape(-1), _np(projmatrix).reshape(-1), _np(campos).reshape(-1), _np(bg)
    radii = np.zeros(P, np.int32)
    xy = np.zeros((P, 2), np.float32)
    depths = np.zeros(P, np.float32)
    cov3Ds = np.zeros((P, 6), np.float32)
    rgb = np.zeros((P, 3), np.float32)
    conic_opacity = np.zeros((P, 4), np.float32)
    tiles = np.zeros(P, np.uint32)
    clamped = np.zeros((P, 3), np.uint8)
    L.oracle_preprocess(ctypes.c_int(P), ctypes.c_int(int(sh_degree)), ctypes.c_int(M), _p(means), _p(scales_),
                        ctypes.c_float(scale_modifier), _p(rots_), _p(opac), _p(shs_), _p(cov_), _p(colors_), _p(view),
                        _p(proj), _p(cam), ctypes.c_int(W), ctypes.c_int(H), ctypes.c_float(tanfovx),
                        ctypes.c_float(tanfovy), _p(radii, ctypes.c_int32), _p(xy), _p(depths), _p(cov3Ds), _p(rgb),
                        _p(conic_opacity), _p(tiles, ctypes.c_uint32), _p(clamped, ctypes.c_uint8))
    offsets = np.zeros(P, np.uint32)
    L.oracle_scan.restype = ctypes.c_int64
    R = int(L.oracle_scan(ctypes.c_int(P), _p(tiles, ctypes.c_uint32), _p(offsets, ctypes.c_uint32)))
    gx, gy = (W + 15) // 16, (H + 15) // 16
    keys_u = np.zeros(max(R, 1), np.uint64)
    vals_u = np.zeros(max(R, 1), np.uint32)
    keys_s = np.zeros(max(R, 1), np.uint64)
    plist = np.zeros(max(R, 1), np.uint32)
    ranges = np.zeros((gx * gy, 2), np.uint32)
    L.oracle_bin_sort(ctypes.c_int(P), ctypes.c_int(R), ctypes.c_int(W), ctypes.c_int(H), _p(xy), _p(depths),
                      _p(radii, ctypes.c_int32), _p(offsets, ctypes.c_uint32), _p(keys_u, ctypes.c_uint64),
                      _p(vals_u, ctypes.c_uint32), _p(keys_s, ctypes.c_uint64), _p(plist, ctypes.c_uint32),
                      _p(ranges, ctypes.c_uint32))
    feats = colors_ if colors_ is not None else rgb
    final_T = np.zeros(W * H, np.float32)
    n_contrib = np.zeros(W * H, np.uint32)
    color = np.zeros((3, H, W), np.float32)
    depth = np.zeros((H, W), np.float32)
    alpha = np.zeros((H, W), np.float32)
    if P > 0:  # rasterize_points.cu:81: P == 0 leaves an all-zero image (no background)
        L.oracle_render_fwd(ctypes.c_int(W), ctypes.c_int(H), _p(ranges, ctypes.c_uint32), _p(plist, ctypes.c_uint32),
                            _p(xy), _p(feats), _p(conic_opacity), _p(depths), _p(bg_), _p(final_T),
                            _p(n_contrib, ctypes.c_uint32), _p(color), _p(depth), _p(alpha))
    return dict(color=color, radii=radii, depth=depth, alpha=alpha, num_rendered=R, xy=xy, depths=depths, cov3D=cov3Ds,
                rgb=rgb, conic_opacity=conic_opacity, tiles_touched=tiles, clamped=clamped, point_offsets=offsets,
                keys_sorted=keys_s[:R], point_list=plist[:R], ranges=ranges, final_T=final_T, n_contrib=n_contrib,
                W=W, H=H, M=M)


def backward(fw: dict, dL_dout_color, *, means3D, viewmatrix, projmatrix, campos, bg, tanfovx, tanfovy, sh_degree=0,
             shs=None, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None, scale_modifier=1.0,
             **_ignored) -> dict:
    """Full backward (Rasterizer::backward, rasterizer_impl.cu:340-434) on the forward dict returned above."""
    L = lib()
    means = _np(means3D)
    P = means.shape[0]
    W, H, M = fw["W"], fw["H"], fw["M"]
    shs_, colors_ = _np(shs), _np(colors_precomp)
    scales_, rots_, cov_ = _np(scales), _np(rotations), _np(cov3D_precomp)
    view, proj, cam, bg_ = _np(viewmatrix).reshape(-1), _np(projmatrix).reshape(-1), _np(campos).reshape(-1), _np(bg)
    dpix = _np(dL_dout_color)
    feats = colors_ if colors_ is not None else fw["rgb"]
    dmean2D = np.zeros((P, 3), np.float32)
    dconic = np.zeros((P, 4), np.float32)
    dopac = np.zeros((P, 1), np.float32)
    dcolors = np.zeros((P, 3), np.float32)
    dmeans = np.zeros((P, 3), np.float32)
    dcov = np.zeros((P, 6), np.float32)
    dsh = np.zeros((P, M, 3), np.float32)
    dscale = np.zeros((P, 3), np.float32)
    drot = np.zeros((P, 4), np.float32)
    if P > 0:
        plist = np.ascontiguousarray(fw["point_list"]) if fw["num_rendered"] > 0 else np.zeros(1, np.uint32)
        L.oracle_render_bwd(ctypes.c_int(W), ctypes.c_int(H), _p(fw["ranges"], ctypes.c_uint32),
                            _p(plist, ctypes.c_uint32), _p(bg_), _p(fw["xy"]), _p(fw["conic_opacity"]), _p(feats),
                            _p(fw["final_T"]), _p(fw["n_contrib"], ctypes.c_uint32), _p(dpix), _p(dmean2D), _p(dconic),
                            _p(dopac), _p(dcolors))
        cov_src = cov_ if cov_ is not None else fw["cov3D"]
        L.oracle_preprocess_bwd(ctypes.c_int(P), ctypes.c_int(int(sh_degree)), ctypes.c_int(M), _p(means),
                                _p(fw["radii"], ctypes.c_int32), _p(shs_), _p(fw["clamped"], ctypes.c_uint8),
                                _p(scales_), _p(rots_), ctypes.c_float(scale_modifier), _p(cov_src), _p(view), _p(proj),
                                ctypes.c_int(W), ctypes.c_int(H), ctypes.c_float(tanfovx), ctypes.c_float(tanfovy),
                                _p(cam), _p(dmean2D), _p(dconic), _p(dcolors), _p(dmeans), _p(dcov),
                                _p(dsh) if M > 0 else None, _p(dscale), _p(drot))
    return dict(dL_dmeans2D=dmean2D, dL_dcolors=dcolors, dL_dopacity=dopac, dL_dmeans3D=dmeans, dL_dcov3D=dcov,
                dL_dsh=dsh, dL_dscales=dscale, dL_drotations=drot, dL_dconic=dconic)


def mark_visible(means3D, viewmatrix, projmatrix):
    L = lib()
    means = _np(means3D)
    P = 0 if means is None else means.shape[0]
    out = np.zeros(P, np.uint8)
    if P:
        L.oracle_mark_visible(ctypes.c_int(P), _p(means), _p(_np(viewmatrix).reshape(-1)),
                              _p(_np(projmatrix).reshape(-1)), _p(out, ctypes.c_uint8))
    return out.astype(bool)


def knn_dist2(points):
    """Brute force O(P^2) — small cases only (use das3r_b200.synth.knn_dist2_cpu, a k-d tree, for large P)."""
    L = lib()
    pts = _np(points)
    P = 0 if pts is None else pts.shape[0]
    out = np.zeros(P, np.float32)
    if P:
        L.oracle_knn(ctypes.c_int(P), _p(pts), _p(out))
    return out
