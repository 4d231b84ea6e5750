"""das3r_frontend_forward_ex / das3r_frontend_backward_ex: active SH degree below 3, render()'s `filtering` selection
(gaussian_renderer/__init__.py:38-46,:85-86,:95,:107,:126), Gaussian ranges and slice-local gradient windows. The full
path (degree 3, no filter, one call) is pinned to the reference's golden vectors in tests/test_gpu_frontend.py; here the
variants are checked against it and against the reference's torch formulation through autograd."""
import numpy as np
import pytest
import torch

import _harness as Hn
from das3r_b200 import load_native

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
EI = torch.empty(0, dtype=torch.int64)
EU = torch.empty(0, dtype=torch.uint8)


def _model(P, seed):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    m = dict(xyz=r(P, 3), rotation=r(P, 4), opacity=r(P, 1), scaling=-3 + 0.5 * r(P, 3), features_dc=r(P, 1, 3),
             features_rest=r(P, 15, 3), conf_static=torch.rand(P, generator=g))
    return {k: v.to(DEV).contiguous() for k, v in m.items()}, g


def _poses(V, g):
    q = torch.tensor([1.0, 0, 0, 0]) + 0.2 * torch.randn(V, 4, generator=g)
    return torch.cat([q, 0.3 * torch.randn(V, 3, generator=g)], dim=1).to(DEV).contiguous()


def _fwd(C, m, poses, D=3, fi=EI):
    return C.frontend_forward_ex(m["xyz"], m["rotation"], m["opacity"], m["scaling"], m["features_dc"], m["features_rest"],
                                 m["conf_static"], EI, poses, D, fi)


def _gin(rows, V, K, g):
    r = lambda *s: torch.randn(*s, generator=g).to(DEV)
    return dict(dop=r(rows, 1), dsc=r(rows, 3), dsh=r(rows, K, 3), dm=[r(rows, 3) for _ in range(V)],
                dr=[r(rows, 4) for _ in range(V)])


def _bwd(C, m, poses, gi, D, fi=EI, outs=(), scratch=EU, first=0, count=-1, finish=True):
    return C.frontend_backward_ex(m["xyz"], m["rotation"], m["opacity"], m["scaling"], m["conf_static"], EI, poses, gi["dop"],
                                  gi["dsc"], gi["dsh"], gi["dm"], gi["dr"], D, fi, list(outs), scratch, first, count, finish)


@pytest.mark.parametrize("D", [0, 1, 2])
def test_active_degree_copies_only_the_active_bands(D):
    C = load_native()
    m, g = _model(3001, 1)
    poses = _poses(3, g)
    full = _fwd(C, m, poses, 3)
    part = _fwd(C, m, poses, D)
    K = (D + 1) ** 2
    assert torch.equal(part[2][:, :K], full[2][:, :K])  # the active coefficients, bit-identical
    assert torch.equal(part[0], full[0]) and torch.equal(part[1], full[1])
    for v in range(3):
        assert torch.equal(part[3][v], full[3][v]) and torch.equal(part[4][v], full[4][v])
    # backward: incoming rows with K coefficients (trimmed) or 16 (full, inactive bands zero) give the same gradients; the
    # outgoing rest rows are [P,15,3] with zeros in the inactive bands, or trimmed to K-1 coefficients
    P, V = 3001, 3
    gi = _gin(P, V, K, g)
    gi16 = dict(gi)
    gi16["dsh"] = torch.cat([gi["dsh"], torch.zeros(P, 16 - K, 3, device=DEV)], dim=1).contiguous()
    a = _bwd(C, m, poses, gi, D)
    b = _bwd(C, m, poses, gi16, 3)
    torch.cuda.synchronize()
    for i in range(8):
        assert torch.equal(a[i], b[i]) or Hn.rel_err(a[i], b[i]) <= 1e-6, i
    assert float(a[5][:, K - 1:].abs().max()) == 0.0
    o = dict(device=DEV)
    nan = float("nan")
    outs = [torch.full((P, 3), nan, **o), torch.full((P, 4), nan, **o), torch.full((P, 1), nan, **o),
            torch.full((P, 3), nan, **o), torch.full((P, 1, 3), nan, **o),
            torch.full((P, K - 1, 3), nan, **o) if K > 1 else torch.empty(0, **o), torch.full((P,), nan, **o),
            torch.full((V, 7), nan, **o)]
    _bwd(C, m, poses, gi, D, outs=outs)
    torch.cuda.synchronize()
    for i in (0, 1, 2, 3, 4, 6, 7):
        assert torch.equal(outs[i], a[i]), i
    if K > 1:
        assert torch.equal(outs[5], a[5][:, :K - 1])


def test_filtering_selects_model_rows():
    """forward: the outputs equal the unfiltered outputs of the selected rows; backward: gradients land on the selected
    rows, the others stay zero; both against torch autograd on the reference's formulation."""
    C = load_native()
    P, V = 2500, 2
    m, g = _model(P, 2)
    poses = _poses(V, g)
    mask = torch.rand(P, generator=g) < 0.6
    fi = mask.nonzero().flatten().to(DEV).contiguous()
    rows = fi.numel()
    mf = dict(m)
    mf["conf_static"] = m["conf_static"][:rows].contiguous()  # indexed by selected row, like the reference's product
    got = _fwd(C, mf, poses, 3, fi)
    sub = {k: (v[fi].contiguous() if k != "conf_static" else mf["conf_static"]) for k, v in m.items()}
    want = _fwd(C, sub, poses, 3)
    for i in range(3):
        assert torch.equal(got[i], want[i])
    for v in range(V):
        assert torch.equal(got[3][v], want[3][v]) and torch.equal(got[4][v], want[4][v])
    gi = _gin(rows, V, 16, g)
    a = _bwd(C, mf, poses, gi, 3, fi)
    b = _bwd(C, sub, poses, gi, 3)
    torch.cuda.synchronize()
    sel = fi
    for i in range(6):
        assert torch.equal(a[i][sel], b[i]), i
        rest = torch.ones(P, dtype=torch.bool, device=DEV)
        rest[sel] = False
        assert float(a[i][rest].abs().max()) == 0.0, i
    assert torch.equal(a[6], b[6]) and Hn.rel_err(a[7], b[7]) <= 1e-6
    # the Python surface takes the boolean mask itself
    from das3r_b200.frontend import render_inputs
    leaves = {k: v.clone().requires_grad_(True) for k, v in mf.items()}
    pz = poses.clone().requires_grad_(True)
    fe = render_inputs(leaves["xyz"], leaves["rotation"], leaves["opacity"], leaves["scaling"], leaves["features_dc"],
                       leaves["features_rest"], leaves["conf_static"], None, pz, sh_degree=3, filtering=mask.to(DEV))
    assert fe.means3D[0].shape == (rows, 3) and torch.equal(fe.opacity, want[0])
    loss = sum((x * y).sum() for x, y in zip(fe.means3D, gi["dm"])) + (fe.opacity * gi["dop"]).sum() + \
        (fe.shs * gi["dsh"]).sum() + (fe.scales * gi["dsc"]).sum() + sum((x * y).sum() for x, y in zip(fe.rotations, gi["dr"]))
    loss.backward()
    assert torch.equal(leaves["xyz"].grad, a[0]) and torch.equal(leaves["features_rest"].grad, a[5])
    assert Hn.rel_err(pz.grad, a[7]) <= 1e-6


@pytest.mark.parametrize("D", [3, 0])
def test_ranges_and_slice_local_windows(D):
    """The backward run range by range into the windows of a slice-major bucket equals the one-call backward; the pose
    gradients are reduced once, after the last range."""
    from das3r_b200 import sharding

    C = load_native()
    P, V = 5003, 3
    K = (D + 1) ** 2
    m, g = _model(P, 3)
    poses = _poses(V, g)
    gi = _gin(P, V, K, g)
    want = _bwd(C, m, poses, gi, D)
    widths = sharding.raw_param_widths(D)
    b = sharding.SlicedGradBucket(P, widths, DEV, 3, tail={"poses": (V, 7)})
    b.flat.fill_(float("nan"))
    scratch = torch.empty(((P + 255) // 256) * V * 16 * 4 + 256, dtype=torch.uint8, device=DEV)
    rest_w = widths.get("features_rest", 0) // 3
    last = len(b.slices) - 1
    for c, (first, count) in enumerate(b.slices):
        w = b.windows(c)
        rest = w["features_rest"].view(count, rest_w, 3) if rest_w else torch.empty(0, device=DEV)
        _bwd(C, m, poses, gi, D, outs=[w["xyz"], w["rotation"], w["opacity"], w["scaling"],
                                       w["features_dc"].view(count, 1, 3), rest, w["conf"].view(count), b.tail("poses")],
             scratch=scratch, first=first, count=count, finish=(c == last))
    torch.cuda.synchronize()
    for i, name in enumerate(("xyz", "rotation", "opacity", "scaling", "features_dc")):
        assert torch.equal(b.gather(name).view(want[i].shape), want[i]), name
    if rest_w:
        assert torch.equal(b.gather("features_rest").view(P, rest_w, 3), want[5][:, :rest_w])
    assert torch.equal(b.gather("conf").view(-1), want[6])
    assert Hn.rel_err(b.tail("poses"), want[7]) <= 1e-6
