"""GPU parity tests (run on the B200 box with `-m gpu`): the new CUDA path, called through the drop-in surface and
directly through the C ABI, against (a) the reference CUDA rasterizer built into oracle/_ref and (b) the C oracle.

Bar (BASELINE.json north_star): integers bit-exact (radii, tiles touched, sorted keys, point list, tile ranges,
n_contrib); forward colour/depth/alpha <= 1e-5 max-abs; gradients <= 1e-3 relative.
"""
import ctypes

import numpy as np
import pytest
import torch

import _harness as Hn
import oracle
from cases import CASE_NAMES, make_case

pytestmark = pytest.mark.gpu

INT_KEYS = ("radii_mismatch", "tiles_mismatch", "xy_mismatch", "depth_mismatch", "clamped_mismatch", "ranges_mismatch",
            "point_list_mismatch", "keys_mismatch", "n_contrib_mismatch")


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def assert_forward(cmp: dict, what: str, exact_floats: bool):
    for k in INT_KEYS:
        if k in cmp:
            assert cmp[k] == 0, f"{what}: {k} = {cmp[k]}"
    if "num_rendered" in cmp:
        assert cmp["num_rendered"][0] == cmp["num_rendered"][1], what
    assert cmp["color_maxabs"] <= Hn.FWD_TOL, f"{what}: colour {cmp['color_maxabs']}"
    if "final_T_maxabs" in cmp:
        assert cmp["final_T_maxabs"] <= Hn.FWD_TOL, what
        assert cmp["rgb_maxabs"] <= Hn.FWD_TOL, what
    if exact_floats and "conic_mismatch" in cmp:
        assert cmp["conic_mismatch"] == 0, f"{what}: conic differs in {cmp['conic_mismatch']} entries"


def assert_grads(cmp: dict, what: str):
    assert cmp, what
    for k, v in cmp.items():
        assert v <= Hn.GRAD_RTOL, f"{what}: {k} rel err {v}"


@pytest.mark.parametrize("name", CASE_NAMES)
def test_case_against_reference_cuda(name, dev):
    if not oracle.reference_available():
        pytest.skip("oracle/_ref not built")
    vi, dL = make_case(name)
    vd = Hn.to_dev(vi, dev)
    ours = Hn.run_ours(vd, dL.to(dev), debug=True)
    ref = Hn.run_ref(vd, dL.to(dev))
    P = vi["means3D"].shape[0]
    assert_forward(Hn.compare_forward(ours, ref, P, has_sh=vi.get("shs") is not None), name, exact_floats=True)
    assert_grads(Hn.compare_grads(ours["grads"], ref["grads"]), name)


@pytest.mark.parametrize("name", CASE_NAMES)
def test_case_against_c_oracle(name, dev):
    vi, dL = make_case(name)
    ours = Hn.run_ours(Hn.to_dev(vi, dev), dL.to(dev), debug=True)
    orc = Hn.run_oracle(vi, dL)
    P = vi["means3D"].shape[0]
    assert_forward(Hn.compare_forward(ours, orc, P, has_sh=vi.get("shs") is not None), name, exact_floats=False)
    assert_grads(Hn.compare_grads(ours["grads"], orc["grads"]), name)
    # additive surface: depth and alpha
    assert float(np.abs(Hn.np_(ours["depth"]) - orc["depth"]).max()) <= Hn.FWD_TOL
    assert float(np.abs(Hn.np_(ours["alpha"]) - orc["alpha"]).max()) <= Hn.FWD_TOL


def test_debug_flag_does_not_change_results(dev):
    vi, dL = make_case("sh3_small")
    vd = Hn.to_dev(vi, dev)
    a = Hn.run_ours(vd, dL.to(dev), debug=True, extras=False)
    b = Hn.run_ours(vd, dL.to(dev), debug=False, extras=False)
    assert torch.equal(a["color"], b["color"]) and torch.equal(a["radii"], b["radii"])
    assert torch.equal(a["point_list"], b["point_list"]) and torch.equal(a["n_contrib"], b["n_contrib"])
    for k in a["grads"]:
        if a["grads"][k] is not None:  # two runs differ only by the order of the atomic additions
            assert Hn.rel_err(a["grads"][k], b["grads"][k]) <= 1e-4


def test_empty_and_single_gaussian(dev):
    import das3r_b200.rasterizer as R

    vi, _ = make_case("sh1_bg_odd")
    vd = Hn.to_dev(vi, dev)
    for n in (0, 1):
        sub = dict(vd)
        for k in ("means3D", "opacities", "shs", "scales", "rotations"):
            sub[k] = vd[k][:n].contiguous()
        ours = Hn.run_ours(sub, torch.ones(3, vi["image_height"], vi["image_width"], device=dev), debug=True)
        cpu = {k: (v.cpu() if isinstance(v, torch.Tensor) else v) for k, v in sub.items()}
        orc = oracle.forward(**cpu)
        assert float(np.abs(Hn.np_(ours["color"]) - orc["color"]).max()) <= Hn.FWD_TOL
        assert ours["radii"].shape == (n,)
        if n == 0:
            assert float(ours["color"].abs().max()) == 0.0  # no background without Gaussians (rasterize_points.cu:81)
            assert ours["grads"]["dL_dmeans3D"].shape == (0, 3)
    assert R.GaussianRasterizer is not None


def test_noncontiguous_inputs_and_side_stream(dev):
    """DAS3R hands over a transposed-slice view of means3D (gaussian_renderer/__init__.py:90); torch streams other than
    the default one must be honoured."""
    vi, dL = make_case("sh3_small")
    vd = Hn.to_dev(vi, dev)
    base = Hn.run_ours(vd, dL.to(dev), debug=False, extras=False)
    m4 = torch.cat([vd["means3D"], torch.ones(vd["means3D"].shape[0], 1, device=dev)], dim=1)
    nc = dict(vd)
    nc["means3D"] = m4.T.contiguous().T[:, :3]
    assert not nc["means3D"].is_contiguous()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        other = Hn.run_ours(nc, dL.to(dev), debug=False, extras=False)
    s.synchronize()
    assert torch.equal(base["color"], other["color"]) and torch.equal(base["radii"], other["radii"])


def test_sh_rows_shorter_than_16(dev):
    """max_sh_degree 1 (M = 4): the generic, unstaged SH path."""
    vi, dL = make_case("sh1_bg_odd")
    vi["shs"] = vi["shs"][:, :4, :].contiguous()
    ours = Hn.run_ours(Hn.to_dev(vi, dev), dL.to(dev), debug=True)
    orc = Hn.run_oracle(vi, dL)
    assert_forward(Hn.compare_forward(ours, orc, vi["means3D"].shape[0]), "M=4", exact_floats=False)
    assert_grads(Hn.compare_grads(ours["grads"], orc["grads"]), "M=4")
    assert ours["grads"]["dL_dsh"].shape == (vi["means3D"].shape[0], 4, 3)


def _crowded_case(n, W=64, H=64, same_depth=False, seed=0):
    """n Gaussians piled into the middle of a small image: one tile list with ~n entries."""
    g = torch.Generator().manual_seed(seed)
    z = torch.full((n,), 2.0) if same_depth else 1.5 + torch.rand(n, generator=g)
    xy = (torch.rand(n, 2, generator=g) - 0.5) * 0.02
    means = torch.stack([xy[:, 0] * z, xy[:, 1] * z, z], dim=1).contiguous()
    vi, _ = make_case("sh3_small")
    vi = dict(vi)
    # anisotropic scales and generic rotations: with isotropic Gaussians the rotation gradient is pure cancellation noise
    rot = torch.randn(n, 4, generator=g)
    rot = (rot / rot.norm(dim=1, keepdim=True)).contiguous()
    vi.update(means3D=means, opacities=torch.full((n, 1), 0.01),
              scales=(0.002 * torch.tensor([1.0, 1.6, 0.6])).repeat(n, 1).contiguous(), rotations=rot,
              shs=(torch.rand(n, 16, 3, generator=g) - 0.5).contiguous(), image_width=W, image_height=H)
    dL = torch.randn(3, H, W, generator=g) / (W * H)
    return vi, dL


@pytest.mark.parametrize("n,same_depth", [(3000, False), (6000, False), (20000, False), (3000, True), (7000, True)])
def test_crowded_tiles_and_equal_depths(n, same_depth, dev):
    """Exercises every tile-sort size class: bucket path (<= 4096), large class (<= 16384), library fallback
    (> 16384), and - with identical depths - the ties-in-index-order rule through the bitonic fallback."""
    if not oracle.reference_available():
        pytest.skip("oracle/_ref not built")
    vi, dL = _crowded_case(n, same_depth=same_depth, seed=n)
    vd = Hn.to_dev(vi, dev)
    ours = Hn.run_ours(vd, dL.to(dev), debug=True)
    ref = Hn.run_ref(vd, dL.to(dev))
    rng = Hn.np_(ours["ranges"])
    assert int((rng[:, 1] - rng[:, 0]).max()) >= n // 2
    assert_forward(Hn.compare_forward(ours, ref, n), f"crowded {n}", exact_floats=True)
    assert_grads(Hn.compare_grads(ours["grads"], ref["grads"]), f"crowded {n}")


def test_mark_visible(dev):
    import das3r_b200.rasterizer as R

    vi, _ = make_case("behind_and_big")
    vd = Hn.to_dev(vi, dev)
    rast = R.GaussianRasterizer(Hn._settings(R, vd))
    got = rast.markVisible(vd["means3D"])
    want = oracle.mark_visible(vi["means3D"], vi["viewmatrix"], vi["projmatrix"])
    assert got.dtype == torch.bool and np.array_equal(got.cpu().numpy(), want)
    assert rast.markVisible(vd["means3D"][:0]).shape == (0,)
    if oracle.reference_available():
        C, _ = oracle.load_reference()
        assert torch.equal(got, C.mark_visible(vd["means3D"], vd["viewmatrix"], vd["projmatrix"]))


@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 33, 1000, 4097, 100_000])
def test_distcuda2(n, dev):
    from simple_knn._C import distCUDA2
    from das3r_b200.synth import knn_dist2_cpu

    g = torch.Generator().manual_seed(n)
    pts = torch.randn(n, 3, generator=g)
    if n >= 33:
        pts[7] = pts[3]  # duplicates count as distance 0 (only the index itself is skipped)
        pts[:20, 2] = 0.5  # a coplanar clump
    got = distCUDA2(pts.to(dev)).cpu()
    if n >= 4:
        want = knn_dist2_cpu(pts)
        assert torch.allclose(got, want, rtol=1e-5, atol=0), float((got - want).abs().max())
    else:
        assert got.shape == (n,)  # fewer than 3 neighbours: the reference leaves FLT_MAX terms in the mean
        assert bool(torch.all(got > 1e30)) or n == 0
    if oracle.reference_available() and n >= 4:
        _, K = oracle.load_reference()
        ref = K.distCUDA2(pts.to(dev)).cpu()
        assert torch.allclose(got, ref, rtol=1e-6, atol=0)


def test_accumulate_mode_and_optional_outputs(dev):
    """backward_into(accumulate=True) adds into caller buffers and leaves unrendered Gaussians untouched."""
    from das3r_b200 import load_native

    C = load_native()
    vi, dL = make_case("behind_and_big")
    vd = Hn.to_dev(vi, dev)
    e = torch.Tensor([])
    args = (vd["bg"], vd["means3D"], e, vd["opacities"], vd["scales"], vd["rotations"], 1.0, e, vd["viewmatrix"],
            vd["projmatrix"], vd["tanfovx"], vd["tanfovy"], vd["image_height"], vd["image_width"], vd["shs"],
            vd["sh_degree"], vd["campos"], False, False)
    R, color, radii, geom, binning, img = C.rasterize_gaussians(*args)
    bargs = (vd["bg"], vd["means3D"], radii, e, vd["scales"], vd["rotations"], 1.0, e, vd["viewmatrix"],
             vd["projmatrix"], vd["tanfovx"], vd["tanfovy"], dL.to(dev), vd["shs"], vd["sh_degree"], vd["campos"], geom,
             R, binning, img, False)
    fresh = C.rasterize_gaussians_backward(*bargs)
    P = vi["means3D"].shape[0]
    def bufs(fill):
        return [e, e, torch.full((P, 1), fill, device=dev), torch.full((P, 3), fill, device=dev), e,
                torch.full((P, 16, 3), fill, device=dev), torch.full((P, 3), fill, device=dev),
                torch.full((P, 4), fill, device=dev)]

    outs = bufs(0.0)
    C.rasterize_gaussians_backward_into(*bargs, outs, True)
    C.rasterize_gaussians_backward_into(*bargs, outs, True)
    for i in (2, 3, 5, 6, 7):
        assert Hn.rel_err(outs[i], 2.0 * fresh[i]) <= 1e-5, i
    marked = bufs(7.0)
    C.rasterize_gaussians_backward_into(*bargs, marked, True)
    invisible = (radii == 0)
    assert bool(invisible.any())
    for i in (2, 3, 5, 6, 7):
        assert bool(torch.all(marked[i][invisible] == 7.0)), i   # unrendered Gaussians are left untouched
    over = bufs(7.0)
    C.rasterize_gaussians_backward_into(*bargs, over, False)     # overwrite mode: same as the fresh tensors
    for i in (2, 3, 5, 6, 7):
        assert torch.equal(over[i].view(-1), fresh[i].view(-1)) or Hn.rel_err(over[i], fresh[i]) <= 1e-5, i


class _Params(ctypes.Structure):
    _fields_ = [("P", ctypes.c_int32), ("D", ctypes.c_int32), ("M", ctypes.c_int32), ("width", ctypes.c_int32),
                ("height", ctypes.c_int32), ("tan_fovx", ctypes.c_float), ("tan_fovy", ctypes.c_float),
                ("scale_modifier", ctypes.c_float), ("prefiltered", ctypes.c_int32), ("debug", ctypes.c_int32),
                ("background", ctypes.c_void_p), ("viewmatrix", ctypes.c_void_p), ("projmatrix", ctypes.c_void_p),
                ("campos", ctypes.c_void_p)]


class _Inputs(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations",
                                               "cov3D_precomp")]


class _Outputs(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in ("out_color", "radii", "out_depth", "out_alpha")]


class _State(ctypes.Structure):
    _fields_ = [("geom", ctypes.c_void_p), ("geom_bytes", ctypes.c_size_t), ("binning", ctypes.c_void_p),
                ("binning_bytes", ctypes.c_size_t), ("image", ctypes.c_void_p), ("image_bytes", ctypes.c_size_t),
                ("num_rendered", ctypes.c_int32)]


class _Grads(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D",
                                               "dL_dsh", "dL_dscales", "dL_drotations", "scratch")] + \
               [("accumulate", ctypes.c_int32)]


def test_c_abi_called_directly_with_ctypes(dev):
    """No torch binding in between: raw device pointers, allocator callbacks, explicit stream."""
    lib = Hn.cabi()
    vi, dL = make_case("sh3_small")
    vd = {k: (v.contiguous() if isinstance(v, torch.Tensor) else v) for k, v in Hn.to_dev(vi, dev).items()}
    P, W, H = vd["means3D"].shape[0], vd["image_width"], vd["image_height"]
    blobs = []
    ALLOC = ctypes.CFUNCTYPE(ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p)

    def alloc(nbytes, _ctx):
        t = torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=dev)
        blobs.append(t)
        return t.data_ptr()

    cb = ALLOC(alloc)
    p = _Params(P, 3, 16, W, H, vd["tanfovx"], vd["tanfovy"], 1.0, 0, 0, vd["bg"].data_ptr(),
                vd["viewmatrix"].data_ptr(), vd["projmatrix"].data_ptr(), vd["campos"].data_ptr())
    inp = _Inputs(vd["means3D"].data_ptr(), vd["shs"].data_ptr(), None, vd["opacities"].data_ptr(),
                  vd["scales"].data_ptr(), vd["rotations"].data_ptr(), None)
    color = torch.empty(3, H, W, device=dev)
    radii = torch.empty(P, dtype=torch.int32, device=dev)
    out = _Outputs(color.data_ptr(), radii.data_ptr(), None, None)
    st = _State()
    stream = torch.cuda.current_stream().cuda_stream
    R = lib.das3r_raster_forward(ctypes.byref(p), ctypes.byref(inp), ctypes.byref(out), cb, None, cb, None, cb, None,
                                 ctypes.byref(st), ctypes.c_void_p(stream))
    assert R >= 0, lib.das3r_last_error()
    assert st.num_rendered == R and len(blobs) == 3
    names = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
             "dL_drotations")
    shapes = ((P, 3), (P, 3), (P, 1), (P, 3), (P, 6), (P, 16, 3), (P, 3), (P, 4))
    gt = {n: torch.empty(*s, device=dev) for n, s in zip(names, shapes)}
    scratch = torch.empty(int(lib.das3r_raster_backward_scratch_bytes(ctypes.c_int32(P))), dtype=torch.uint8, device=dev)
    g = _Grads(*[gt[n].data_ptr() for n in names], scratch.data_ptr(), 0)
    dLd = dL.to(dev).contiguous()
    rc = lib.das3r_raster_backward(ctypes.byref(p), ctypes.byref(inp), ctypes.c_void_p(radii.data_ptr()),
                                   ctypes.byref(st), ctypes.c_void_p(dLd.data_ptr()), ctypes.byref(g),
                                   ctypes.c_void_p(stream))
    assert rc == 0, lib.das3r_last_error()
    torch.cuda.synchronize()
    ours = Hn.run_ours(Hn.to_dev(vi, dev), dL.to(dev), debug=False, extras=False)
    assert torch.equal(color, ours["color"]) and torch.equal(radii, ours["radii"]) and R == ours["num_rendered"]
    assert Hn.rel_err(gt["dL_dmeans3D"], ours["grads"]["dL_dmeans3D"]) <= 1e-5
    assert Hn.rel_err(gt["dL_dsh"], ours["grads"]["dL_dsh"]) <= 1e-5
    orc = Hn.run_oracle(vi, dL)
    assert float(np.abs(color.cpu().numpy() - orc["color"]).max()) <= Hn.FWD_TOL
    # error path: exactly one of shs / colors_precomp
    bad = _Inputs(vd["means3D"].data_ptr(), None, None, vd["opacities"].data_ptr(), vd["scales"].data_ptr(),
                  vd["rotations"].data_ptr(), None)
    rc = lib.das3r_raster_forward(ctypes.byref(p), ctypes.byref(bad), ctypes.byref(out), cb, None, cb, None, cb, None,
                                  ctypes.byref(st), ctypes.c_void_p(stream))
    assert rc == -1 and b"exactly one" in lib.das3r_last_error()


def test_full_size_c1_and_c2_against_reference(dev):
    """BASELINE.json configs at full size: C1 (100k) against reference CUDA and the C oracle, C2 (1M Gaussians,
    512x288) against the reference CUDA, plus size-independent properties."""
    from das3r_b200.synth import CONFIGS, grad_seed, synth_scene, view_inputs

    for name, view in (("C1", 0), ("C2", 7)):
        sc = synth_scene(**CONFIGS[name])
        vi = view_inputs(sc, view)
        vi["means3D"] = vi["means3D"].contiguous()
        dL = grad_seed(sc)
        vd = Hn.to_dev(vi, dev)
        ours = Hn.run_ours(vd, dL.to(dev), debug=True)
        P = sc.P
        # properties that need no oracle
        keys = ours["keys_sorted"]
        assert bool(torch.all(keys[1:] >= keys[:-1]))
        assert int(ours["tiles_touched"].sum()) == ours["num_rendered"]
        rng = ours["ranges"].long()
        assert int((rng[:, 1] - rng[:, 0]).sum()) == ours["num_rendered"]
        assert bool(torch.all((ours["radii"] > 0) == (ours["tiles_touched"] > 0)))
        assert bool(torch.isfinite(ours["color"]).all())
        for g in ours["grads"].values():
            if g is not None:
                assert bool(torch.isfinite(g).all())
        # linearity of the backward in dL (the backward is a vector-Jacobian product)
        twice = Hn.run_ours(vd, 2.0 * dL.to(dev), debug=False, extras=False)
        assert Hn.rel_err(twice["grads"]["dL_dmeans3D"], 2.0 * ours["grads"]["dL_dmeans3D"]) <= 1e-4
        if oracle.reference_available():
            ref = Hn.run_ref(vd, dL.to(dev))
            assert_forward(Hn.compare_forward(ours, ref, P), name, exact_floats=True)
            assert_grads(Hn.compare_grads(ours["grads"], ref["grads"]), name)
        if name == "C1":
            orc = Hn.run_oracle(vi, dL)
            assert_forward(Hn.compare_forward(ours, orc, P), name + " oracle", exact_floats=False)
            assert_grads(Hn.compare_grads(ours["grads"], orc["grads"]), name + " oracle")
        del ours, twice
        torch.cuda.empty_cache()
