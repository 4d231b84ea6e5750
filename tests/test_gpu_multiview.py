"""das3r_raster_backward_blend + das3r_raster_backward_finish: the backward split in two halves so that the views of a
training step share one pass over the Gaussians (no counterpart in the reference; its backward handles one view per
call, rasterizer_impl.cu:340-434, and autograd sums).

Bar: the sum the fused pass leaves equals the sum of the per-view gradients of the regular backward (which the other
parity tests pin to the reference) within the gradient tolerance - run-to-run atomic order and the order of the float
additions are the only differences. Covered: the fused path (degree 3, shared SH rows), the per-view fallback (degree 1),
accumulate on / off, one view, more views than one launch takes, Gaussians invisible in some or all views.
"""
import pytest
import torch

import _harness as Hn
from das3r_b200 import load_native
from das3r_b200.synth import synth_scene, view_inputs

pytestmark = pytest.mark.gpu

NAMES = ("dL_dopacity", "dL_dmeans3D", "dL_dsh", "dL_dscales", "dL_drotations")
IDX = (2, 3, 5, 6, 7)  # their position in the 8-tuple of rasterize_gaussians_backward


def _scene_views(P, V, W, H, D, seed, dev):
    sc = synth_scene(P=P, V=max(V, 2), W=W, H=H, seed=seed, sh_degree=D)
    views = []
    for v in range(V):
        vi = view_inputs(sc, v, sh_degree=D)
        views.append({k: (x.contiguous().to(dev) if isinstance(x, torch.Tensor) else x) for k, x in vi.items()})
    shs = views[0]["shs"]
    for vi in views:  # one SH array for all views, as in DAS3R (the pose moves means and rotations only)
        vi["shs"] = shs
    g = torch.Generator().manual_seed(seed + 100)
    dLs = [(torch.randn(3, H, W, generator=g) / (W * H)).to(dev) for _ in range(V)]
    return views, dLs


def _forward(C, vi, D):
    e = torch.Tensor([])
    return C.rasterize_gaussians(vi["bg"], vi["means3D"], e, vi["opacities"], vi["scales"], vi["rotations"], 1.0, e,
                                 vi["viewmatrix"], vi["projmatrix"], vi["tanfovx"], vi["tanfovy"], vi["image_height"],
                                 vi["image_width"], vi["shs"], D, vi["campos"], False, False)


def _regular_sum(C, views, dLs, D):
    """Sum over the views of the regular per-view backward (fresh buffers per view, summed in float64)."""
    e = torch.Tensor([])
    tot = None
    states = []
    for vi, dL in zip(views, dLs):
        R, color, radii, geom, binning, img = _forward(C, vi, D)
        states.append((R, radii, geom, binning, img))
        out = C.rasterize_gaussians_backward(vi["bg"], vi["means3D"], radii, e, vi["scales"], vi["rotations"], 1.0, e,
                                             vi["viewmatrix"], vi["projmatrix"], vi["tanfovx"], vi["tanfovy"], dL,
                                             vi["shs"], D, vi["campos"], geom, R, binning, img, False)
        g = [out[i].double() for i in IDX]
        tot = g if tot is None else [a + b for a, b in zip(tot, g)]
    return tot, states


def _fused(C, views, dLs, states, D, outs, accumulate, per_view=None):
    """per_view: optional list (one entry per view) of (dL_dmeans3D, dL_drotations) tensors or None"""
    fv = []
    e = torch.Tensor([])
    for i, (vi, dL, (R, radii, geom, binning, img)) in enumerate(zip(views, dLs, states)):
        acc = C.rasterize_gaussians_backward_blend(vi["bg"], dL, vi["means3D"].shape[0], geom, R, binning, img, False)
        pm, pr = (per_view[i] if per_view and per_view[i] is not None else (e, e))
        fv.append((vi["means3D"], radii, vi["scales"], vi["rotations"], vi["viewmatrix"], vi["projmatrix"],
                   float(vi["tanfovx"]), float(vi["tanfovy"]), vi["campos"], geom, acc, int(vi["image_height"]),
                   int(vi["image_width"]), pm, pr))
    C.rasterize_gaussians_backward_finish(fv, views[0]["shs"], D, 1.0, outs, accumulate, False)
    torch.cuda.synchronize()


def _fresh(P, M, dev, fill=float("nan")):
    shapes = ((P, 1), (P, 3), (P, M, 3), (P, 3), (P, 4))
    return [torch.full(s, fill, device=dev) for s in shapes]


def _check(outs, want, what):
    for n, o, w in zip(NAMES, outs, want):
        assert bool(torch.isfinite(o).all()), f"{what}: {n} has elements the call did not write"
        err = Hn.rel_err(o.double().cpu().numpy(), w.cpu().numpy())
        assert err <= Hn.GRAD_RTOL, f"{what}: {n} differs from the sum of per-view backwards by {err}"


@pytest.mark.parametrize("D,V", [(3, 4), (1, 3), (3, 1), (3, 18), (0, 3), (2, 5), (0, 18)])
def test_sum_over_views_matches_the_per_view_backward(D, V):
    dev = torch.device("cuda:0")
    C = load_native()
    P, W, H = 20_000, 160, 96
    views, dLs = _scene_views(P, V, W, H, D, seed=11 + D, dev=dev)
    M = views[0]["shs"].shape[1]
    want, states = _regular_sum(C, views, dLs, D)

    outs = _fresh(P, M, dev)
    _fused(C, views, dLs, states, D, outs, accumulate=False)
    _check(outs, want, f"D={D} V={V} overwrite")

    # accumulate: a second identical call on top of the first doubles everything, nothing else is touched
    _fused(C, views, dLs, states, D, outs, accumulate=True)
    _check(outs, [2 * w for w in want], f"D={D} V={V} accumulate")


@pytest.mark.parametrize("D,M4", [(3, False), (1, False), (0, False), (1, True)])
def test_per_view_mean_and_rotation_gradients(D, M4):
    """Views 0 and 2 ask for their own dL_dmeans3D / dL_drotations (what DAS3R needs to go back through each view's
    pose); view 1 joins the summed outputs. D=1 exercises the per-view fallback, which supports this only when
    accumulating."""
    dev = torch.device("cuda:0")
    C = load_native()
    V, P, W, H = 3, 8_000, 128, 80
    views, dLs = _scene_views(P, V, W, H, D, seed=21, dev=dev)
    if M4:  # 4-coefficient rows (max_sh_degree 1): not the fused kernel but one accumulating one-view pass per view
        sh4 = views[0]["shs"][:, :4].contiguous()
        for vi in views:
            vi["shs"] = sh4
    M = views[0]["shs"].shape[1]
    e = torch.Tensor([])
    singles, states = [], []
    for vi, dL in zip(views, dLs):
        R, color, radii, geom, binning, img = _forward(C, vi, D)
        states.append((R, radii, geom, binning, img))
        out = C.rasterize_gaussians_backward(vi["bg"], vi["means3D"], radii, e, vi["scales"], vi["rotations"], 1.0, e,
                                             vi["viewmatrix"], vi["projmatrix"], vi["tanfovx"], vi["tanfovy"], dL,
                                             vi["shs"], D, vi["campos"], geom, R, binning, img, False)
        singles.append([out[i].double() for i in IDX])
    accumulate = False  # every path overwrites: fused kernel, or the per-view fallback clearing a view's own outputs
    fill = float("nan")
    outs = _fresh(P, M, dev, fill=fill)
    per_view = [(torch.full((P, 3), fill, device=dev), torch.full((P, 4), fill, device=dev)), None,
                (torch.full((P, 3), fill, device=dev), torch.full((P, 4), fill, device=dev))]
    _fused(C, views, dLs, states, D, outs, accumulate=accumulate, per_view=per_view)
    want = [sum(s[k] for s in singles) for k in range(5)]
    want[1] = singles[1][1]  # summed dL_dmeans3D: view 1 only
    want[4] = singles[1][4]  # summed dL_drotations: view 1 only
    _check(outs, want, "summed outputs")
    for v in (0, 2):
        for got, w, n in ((per_view[v][0], singles[v][1], "dL_dmeans3D"), (per_view[v][1], singles[v][4], "dL_drotations")):
            assert bool(torch.isfinite(got).all())
            err = Hn.rel_err(got.double().cpu().numpy(), w.cpu().numpy())
            assert err <= Hn.GRAD_RTOL, f"view {v} {n}: {err}"


def test_gaussians_outside_every_view_get_zeros_or_stay_untouched():
    dev = torch.device("cuda:0")
    C = load_native()
    D, V, P, W, H = 3, 3, 6_000, 128, 80
    views, dLs = _scene_views(P, V, W, H, D, seed=5, dev=dev)
    hidden = torch.arange(0, P, 7, device=dev)
    for vi in views:  # behind the camera in all views
        m = vi["means3D"].clone()
        m[hidden, 2] = -1.0
        vi["means3D"] = m
    want, states = _regular_sum(C, views, dLs, D)
    for r in states:
        assert int((r[1][hidden] > 0).sum()) == 0
    outs = _fresh(P, 16, dev)
    _fused(C, views, dLs, states, D, outs, accumulate=False)
    _check(outs, want, "overwrite")
    for o in outs:
        assert float(o[hidden].abs().max()) == 0.0
    marked = _fresh(P, 16, dev, fill=3.0)
    _fused(C, views, dLs, states, D, marked, accumulate=True)
    for o in marked:
        assert bool((o[hidden] == 3.0).all()), "accumulate mode wrote to a Gaussian no view rendered"


@pytest.mark.parametrize("D", [3, 1])
def test_multiview_rasterizer_matches_the_per_view_module(D):
    """MultiViewRasterizer (one autograd node for all views, streams + fused gradient pass) against the drop-in
    GaussianRasterizer called once per view: identical images and radii, gradients within tolerance."""
    from das3r_b200.multiview import MultiViewRasterizer
    from das3r_b200.rasterizer import GaussianRasterizationSettings, GaussianRasterizer

    dev = torch.device("cuda:0")
    V, P, W, H = 4, 12_000, 144, 96
    views, dLs = _scene_views(P, V, W, H, D, seed=31, dev=dev)

    def leaves():
        shared = [views[0][k].clone().requires_grad_(True) for k in ("opacities", "shs", "scales")]
        per = [(vi["means3D"].clone().requires_grad_(True), vi["rotations"].clone().requires_grad_(True)) for vi in views]
        return shared, per

    settings = [GaussianRasterizationSettings(
        image_height=H, image_width=W, tanfovx=vi["tanfovx"], tanfovy=vi["tanfovy"], bg=vi["bg"], scale_modifier=1.0,
        viewmatrix=vi["viewmatrix"], projmatrix=vi["projmatrix"], sh_degree=D, campos=vi["campos"], prefiltered=False,
        debug=False) for vi in views]

    (op_a, sh_a, sc_a), per_a = leaves()
    want_colors = []
    loss = 0
    for rs, (m, r), dL in zip(settings, per_a, dLs):
        m2d = torch.zeros(P, 3, device=dev, requires_grad=True)
        color, radii = GaussianRasterizer(rs)(means3D=m, means2D=m2d, opacities=op_a, shs=sh_a, scales=sc_a, rotations=r)
        want_colors.append(color.detach())
        loss = loss + (color * dL).sum()
    loss.backward()

    (op_b, sh_b, sc_b), per_b = leaves()
    colors, radii = MultiViewRasterizer(streams=3)(settings, [m for m, _ in per_b], [r for _, r in per_b], op_b, sh_b,
                                                   sc_b)
    sum((c * dL).sum() for c, dL in zip(colors, dLs)).backward()
    torch.cuda.synchronize()
    for c, w in zip(colors, want_colors):
        assert float((c.detach() - w).abs().max()) == 0.0
    for name, a, b in (("opacities", op_a, op_b), ("shs", sh_a, sh_b), ("scales", sc_a, sc_b)):
        assert Hn.rel_err(b.grad, a.grad) <= Hn.GRAD_RTOL, name
    for v, ((ma, ra), (mb, rb)) in enumerate(zip(per_a, per_b)):
        assert Hn.rel_err(mb.grad, ma.grad) <= Hn.GRAD_RTOL, f"means3D of view {v}"
        assert Hn.rel_err(rb.grad, ra.grad) <= Hn.GRAD_RTOL, f"rotations of view {v}"


@pytest.mark.parametrize("D,V", [(3, 5), (1, 2), (3, 18), (0, 3), (2, 4), (0, 17)])
def test_forward_split_is_bit_identical_to_the_one_call_forward(D, V):
    """das3r_raster_forward_preprocess (all views in one call: one kernel when they share degree-3 SH rows, the one-view
    kernel otherwise, chunks of 16) + das3r_raster_forward_render per view against das3r_raster_forward: image, radii,
    num_rendered and the whole geometry blob byte for byte."""
    dev = torch.device("cuda:0")
    C = load_native()
    P, W, H = 30_000, 176, 112
    views, _ = _scene_views(P, V, W, H, D, seed=40 + D, dev=dev)
    want = [_forward(C, vi, D) for vi in views]
    pre = C.rasterize_gaussians_forward_preprocess(
        [(vi["bg"], vi["means3D"], vi["rotations"], vi["viewmatrix"], vi["projmatrix"], float(vi["tanfovx"]),
          float(vi["tanfovy"]), H, W, vi["campos"]) for vi in views], views[0]["opacities"], views[0]["scales"],
        views[0]["shs"], D, 1.0, False, False)
    for vi, (color, radii, geom, img), w in zip(views, pre, want):
        R, binning = C.rasterize_gaussians_forward_render(vi["bg"], color, radii, geom, img, False)
        torch.cuda.synchronize()
        assert R == w[0]
        assert torch.equal(color, w[1]) and torch.equal(radii, w[2])
        L = Hn.layout(P, W, H, R)
        vis = radii > 0  # records and clamp bits of Gaussians that are not rendered are never written
        rec = lambda g: g[L.g_recA:L.g_recA + 48 * P].view(torch.int32).view(P, 12)[vis]
        assert torch.equal(rec(geom), rec(w[3]))
        tiles = lambda g: g[L.g_tiles_touched:L.g_tiles_touched + 4 * P].view(torch.int32)
        assert torch.equal(tiles(geom), tiles(w[3]))
        clamp = lambda g: g[L.g_clamped:L.g_clamped + P][vis]
        assert torch.equal(clamp(geom), clamp(w[3]))


@pytest.mark.parametrize("D", [3, 0])
def test_more_views_than_one_launch_with_per_view_outputs(D):
    """18 views (one launch takes 16), every view with its own dL_dmeans3D / dL_drotations, overwrite mode: the second
    launch must overwrite ITS views' buffers (NaN-poisoned here) while it adds to the summed outputs."""
    dev = torch.device("cuda:0")
    C = load_native()
    V, P, W, H = 18, 6_000, 112, 80
    views, dLs = _scene_views(P, V, W, H, D, seed=23, dev=dev)
    e = torch.Tensor([])
    singles, states = [], []
    for vi, dL in zip(views, dLs):
        R, color, radii, geom, binning, img = _forward(C, vi, D)
        states.append((R, radii, geom, binning, img))
        out = C.rasterize_gaussians_backward(vi["bg"], vi["means3D"], radii, e, vi["scales"], vi["rotations"], 1.0, e,
                                             vi["viewmatrix"], vi["projmatrix"], vi["tanfovx"], vi["tanfovy"], dL,
                                             vi["shs"], D, vi["campos"], geom, R, binning, img, False)
        singles.append([out[i].double() for i in IDX])
    nan = float("nan")
    outs = _fresh(P, 16, dev)
    per_view = [(torch.full((P, 3), nan, device=dev), torch.full((P, 4), nan, device=dev)) for _ in range(V)]
    _fused(C, views, dLs, states, D, outs, accumulate=False, per_view=per_view)
    for k in (0, 2, 3):  # opacity, sh, scales are summed over all 18 views
        want = sum(s[k] for s in singles)
        assert Hn.rel_err(outs[k].double().cpu().numpy(), want.cpu().numpy()) <= Hn.GRAD_RTOL, NAMES[k]
    for v in range(V):
        for got, w, n in ((per_view[v][0], singles[v][1], "dL_dmeans3D"), (per_view[v][1], singles[v][4], "dL_drotations")):
            assert bool(torch.isfinite(got).all()), f"view {v} {n}: not every element was written"
            assert Hn.rel_err(got.double().cpu().numpy(), w.cpu().numpy()) <= Hn.GRAD_RTOL, f"view {v} {n}"


@pytest.mark.parametrize("D", [0, 1, 2, 3])
def test_trimmed_rows_and_gaussian_ranges(D):
    """das3r_raster_backward_finish_ex: dL_dsh rows trimmed to the (D+1)^2 active coefficients, computed range by range
    (absolute-indexed buffers and slice-local windows), equal the full-row one-call result."""
    dev = torch.device("cuda:0")
    C = load_native()
    V, P, W, H = 3, 5_003, 128, 80
    K = (D + 1) ** 2
    views, dLs = _scene_views(P, V, W, H, D, seed=61 + D, dev=dev)
    want, states = _regular_sum(C, views, dLs, D)
    e = torch.Tensor([])
    fv = []
    for vi, dL, (R, radii, geom, binning, img) in zip(views, dLs, states):
        acc = C.rasterize_gaussians_backward_blend(vi["bg"], dL, P, geom, R, binning, img, False)
        fv.append((vi["means3D"], radii, vi["scales"], vi["rotations"], vi["viewmatrix"], vi["projmatrix"],
                   float(vi["tanfovx"]), float(vi["tanfovy"]), vi["campos"], geom, acc, H, W, e, e))
    nan = float("nan")
    shapes = ((P, 1), (P, 3), (P, K, 3), (P, 3), (P, 4))
    outs = [torch.full(s, nan, device=dev) for s in shapes]
    ranges = [(0, 2048), (2048, 1024), (3072, P - 3072)]
    for first, count in ranges:
        C.rasterize_gaussians_backward_finish_ex(fv, views[0]["shs"], D, 1.0, outs, False, False, K, first, count)
    torch.cuda.synchronize()
    want_t = list(want)
    want_t[2] = want[2][:, :K]
    assert float(want[2][:, K:].abs().max()) == 0.0 if K < 16 else True
    _check(outs, want_t, f"D={D} trimmed, absolute ranges")
    # slice-local windows (what a slice-major gradient bucket hands over)
    for first, count in ranges:
        loc = [torch.full((count,) + s[1:], nan, device=dev) for s in shapes]
        C.rasterize_gaussians_backward_finish_ex(fv, views[0]["shs"], D, 1.0, loc, False, False, K, first, count)
        torch.cuda.synchronize()
        _check(loc, [w[first:first + count] for w in want_t], f"D={D} trimmed, local window at {first}")
    # accumulate on top
    for first, count in ranges:
        C.rasterize_gaussians_backward_finish_ex(fv, views[0]["shs"], D, 1.0, outs, True, False, K, first, count)
    torch.cuda.synchronize()
    _check(outs, [2 * w for w in want_t], f"D={D} trimmed, accumulate")


@pytest.mark.parametrize("D,V", [(3, 5), (0, 3), (1, 18)])
def test_batched_render_and_blend_match_the_per_view_calls(D, V):
    """das3r_raster_forward_render_batch / das3r_raster_backward_blend_batch (one launch per stage for all views)
    against the per-view entry points: images, radii, point lists, n_contrib, final_T bit-identical; the accumulators
    equal up to the order of the atomic additions."""
    dev = torch.device("cuda:0")
    C = load_native()
    P, W, H = 25_000, 176, 112
    views, dLs = _scene_views(P, V, W, H, D, seed=70 + D, dev=dev)

    def pre():
        return C.rasterize_gaussians_forward_preprocess(
            [(vi["bg"], vi["means3D"], vi["rotations"], vi["viewmatrix"], vi["projmatrix"], float(vi["tanfovx"]),
              float(vi["tanfovy"]), H, W, vi["campos"]) for vi in views], views[0]["opacities"], views[0]["scales"],
            views[0]["shs"], D, 1.0, False, True)

    pa = pre()
    single = [C.rasterize_gaussians_forward_render(vi["bg"], p[0], p[1], p[2], p[3], True) for vi, p in zip(views, pa)]
    pb = pre()
    batch = C.rasterize_gaussians_forward_render_batch([(vi["bg"], p[0], p[1], p[2], p[3]) for vi, p in zip(views, pb)], True)
    torch.cuda.synchronize()
    for v in range(V):
        Ra, bina = single[v]
        Rb, binb = batch[v]
        assert Ra == Rb
        assert torch.equal(pa[v][0], pb[v][0]) and torch.equal(pa[v][1], pb[v][1])
        L = Hn.layout(P, W, H, Ra)
        for off, cnt in ((L.b_point_list, 4 * Ra), (L.b_keys_sorted, 8 * Ra)):
            assert torch.equal(bina[off:off + cnt], binb[off:off + cnt])
        for off, cnt in ((L.i_final_T, 4 * W * H), (L.i_n_contrib, 4 * W * H)):
            assert torch.equal(pa[v][3][off:off + cnt], pb[v][3][off:off + cnt])
    acc_single = [C.rasterize_gaussians_backward_blend(vi["bg"], dL, P, pa[v][2], single[v][0], single[v][1], pa[v][3], True)
                  for v, (vi, dL) in enumerate(zip(views, dLs))]
    acc_batch = C.rasterize_gaussians_backward_blend_batch(
        [(vi["bg"], dL, pb[v][2], batch[v][0], batch[v][1], pb[v][3]) for v, (vi, dL) in enumerate(zip(views, dLs))], P, True)
    torch.cuda.synchronize()
    for a, b in zip(acc_single, acc_batch):
        assert Hn.rel_err(b[:, :9], a[:, :9]) <= 1e-4
