"""Several views of one set of Gaussians in one autograd node - the shape of a data-parallel DAS3R step.

DAS3R renders one view per iteration (train_gui.py:555-559) and poses the Gaussians for that view itself
(gaussian_renderer/__init__.py:83-92): every view gets its own `means3D` and `rotations`, while opacities, SH
coefficients and scales are shared. A step that renders n views (this repository's multi-GPU plan: each rank takes n
views of the step) can therefore

  * run the n forwards and the n backward blends side by side on a few CUDA streams, each fed by its own host thread
    (the forward blocks its thread once on the read-back of num_rendered; the binding releases the GIL), and
  * turn the n screen-space gradient buffers into Gaussian gradients in ONE pass (`das3r_raster_backward_finish`) that
    reads the shared SH rows once and writes dL_dsh / dL_dscales / dL_dopacity once, summed over the views, plus every
    view's own dL_dmeans3D / dL_drotations.

This has no counterpart in the reference package; `GaussianRasterizer` (rasterizer.py) remains the drop-in per-view
surface and gives the same numbers (tests/test_gpu_multiview.py). Only the input variant DAS3R trains with is
supported here: SH colours + scale/rotation; use the per-view surface for anything else.
"""
from __future__ import annotations

from concurrent.futures import ThreadPoolExecutor
from typing import Sequence

import torch

from . import load_native
from .rasterizer import GaussianRasterizationSettings

_C = load_native()
_EMPTY = torch.Tensor([])


class _Lanes:
    """S CUDA streams of one device, each with a host thread of its own."""

    def __init__(self, device: torch.device, n: int):
        self.device = device
        self.streams = [torch.cuda.Stream(device=device) for _ in range(n)]
        self.workers = ThreadPoolExecutor(max_workers=n)

    def run(self, n_items: int, fn):
        """fn(i) for i in range(n_items); item i runs on stream i % S, after everything queued on the caller's current
        stream; the caller's stream then waits for all of them. Returns the results in order."""
        main = torch.cuda.current_stream(self.device)
        start = torch.cuda.Event()
        start.record(main)
        S = len(self.streams)
        out = [None] * n_items

        def lane(si):
            torch.cuda.set_device(self.device)
            st = self.streams[si]
            with torch.cuda.stream(st):
                st.wait_event(start)
                for i in range(si, n_items, S):
                    out[i] = fn(i)

        for f in [self.workers.submit(lane, si) for si in range(min(S, n_items))]:
            f.result()
        for st in self.streams[:min(S, n_items)]:
            main.wait_stream(st)
        return out


def _crosses_to(stream, *tensors):
    # allocated on a lane's stream, read later on `stream`: keep the caching allocator from recycling it too early
    for t in tensors:
        if isinstance(t, torch.Tensor) and t.is_cuda:
            t.record_stream(stream)


class _RasterizeViews(torch.autograd.Function):
    @staticmethod
    def forward(ctx, settings, lanes, opacities, shs, scales, *mr):
        n = len(settings)
        means = mr[0::2]
        rots = mr[1::2]
        main = torch.cuda.current_stream(opacities.device)

        # first half for all views at once on the caller's stream: views that share the SH rows are preprocessed by one
        # kernel that reads them once (das3r_raster_forward_preprocess)
        rs0 = settings[0]
        pre = _C.rasterize_gaussians_forward_preprocess(
            [(rs.bg, means[i], rots[i], rs.viewmatrix, rs.projmatrix, float(rs.tanfovx), float(rs.tanfovy),
              int(rs.image_height), int(rs.image_width), rs.campos) for i, rs in enumerate(settings)],
            opacities, scales, shs, rs0.sh_degree, float(rs0.scale_modifier), bool(rs0.prefiltered), bool(rs0.debug))

        def fwd(i):
            # second half per view on its lane: read-back of num_rendered, binning, sort, rect lists, blend
            color, radii, geom, img = pre[i]
            lane_stream = torch.cuda.current_stream(opacities.device)
            for t in (color, radii, geom, img):  # allocated on the caller's stream, used on this lane
                if t.numel():
                    t.record_stream(lane_stream)
            R, binning = _C.rasterize_gaussians_forward_render(settings[i].bg, color, radii, geom, img,
                                                               bool(settings[i].debug))
            return (R, color, radii, geom, binning, img)

        res = lanes.run(n, fwd)
        ctx.settings, ctx.lanes, ctx.n = settings, lanes, n
        ctx.num_rendered = [r[0] for r in res]
        colors = [r[1] for r in res]
        radii = [r[2] for r in res]
        saved = [opacities, shs, scales, *mr]
        for r in res:
            saved += [r[2], r[3], r[4], r[5]]
        ctx.save_for_backward(*saved)
        ctx.mark_non_differentiable(*radii)
        return (*colors, *radii)

    @staticmethod
    def backward(ctx, *grads):
        n, settings, lanes = ctx.n, ctx.settings, ctx.lanes
        sv = ctx.saved_tensors
        opacities, shs, scales = sv[0], sv[1], sv[2]
        mr = sv[3:3 + 2 * n]
        st = sv[3 + 2 * n:]
        dev = opacities.device
        main = torch.cuda.current_stream(dev)
        P = opacities.shape[0]
        M = shs.shape[1]

        def blend(i):
            rs = settings[i]
            radii, geom, binning, img = st[4 * i:4 * i + 4]
            g = grads[i]
            if g is None:
                g = torch.zeros(3, rs.image_height, rs.image_width, device=dev)
            acc = _C.rasterize_gaussians_backward_blend(rs.bg, g.contiguous(), P, geom, ctx.num_rendered[i], binning,
                                                        img, rs.debug)
            _crosses_to(main, acc)
            return acc

        accs = lanes.run(n, blend)
        fused = all(rs.sh_degree == 3 for rs in settings) and M == 16
        new = torch.empty if fused else torch.zeros  # the per-view fallback can only add
        o = dict(device=dev, dtype=torch.float32)
        d_opac, d_shs, d_scales = new(P, 1, **o), new(P, M, 3, **o), new(P, 3, **o)
        d_means_sum, d_rots_sum = new(P, 3, **o), new(P, 4, **o)  # nothing is summed into them: every view has its own
        d_mr = []
        views = []
        for i, rs in enumerate(settings):
            radii, geom, binning, img = st[4 * i:4 * i + 4]
            dm, dr = new(P, 3, **o), new(P, 4, **o)
            d_mr += [dm, dr]
            views.append((mr[2 * i], radii, scales, mr[2 * i + 1], rs.viewmatrix, rs.projmatrix, float(rs.tanfovx),
                          float(rs.tanfovy), rs.campos, geom, accs[i], int(rs.image_height), int(rs.image_width), dm,
                          dr))
        _C.rasterize_gaussians_backward_finish(views, shs, settings[0].sh_degree, float(settings[0].scale_modifier),
                                               [d_opac, d_means_sum, d_shs, d_scales, d_rots_sum], not fused,
                                               bool(settings[0].debug))
        return (None, None, d_opac, d_shs, d_scales, *d_mr)


class MultiViewRasterizer:
    """`MultiViewRasterizer(streams=3)(settings, means3D, rotations, opacities, shs, scales)` -> (colors, radii), two
    lists with one entry per view; differentiable w.r.t. opacities, shs, scales and every view's means3D / rotations.

    settings   sequence of GaussianRasterizationSettings, one per view (same sh_degree and scale_modifier)
    means3D    sequence of [P,3] tensors, the Gaussians as posed for each view
    rotations  sequence of [P,4] tensors, likewise
    opacities [P,1], shs [P,M,3], scales [P,3]   shared by all views
    """

    def __init__(self, streams: int = 3):
        self.n_streams = max(1, int(streams))
        self._lanes = {}

    def _lanes_for(self, device):
        key = (device.type, device.index if device.index is not None else torch.cuda.current_device())
        if key not in self._lanes:
            self._lanes[key] = _Lanes(torch.device(*key), self.n_streams)
        return self._lanes[key]

    def __call__(self, settings: Sequence[GaussianRasterizationSettings], means3D: Sequence[torch.Tensor],
                 rotations: Sequence[torch.Tensor], opacities: torch.Tensor, shs: torch.Tensor, scales: torch.Tensor):
        n = len(settings)
        if n == 0 or len(means3D) != n or len(rotations) != n:
            raise ValueError("settings, means3D and rotations need one entry per view")
        if not opacities.is_cuda:
            raise RuntimeError("opacities must be a CUDA tensor: this rasterizer has no CPU path")
        if shs is None or shs.dim() != 3 or scales is None:
            raise Exception("MultiViewRasterizer takes SH colours and scale/rotation inputs; use GaussianRasterizer "
                            "for precomputed colours or covariances")
        if any(rs.sh_degree != settings[0].sh_degree or rs.scale_modifier != settings[0].scale_modifier or
               rs.prefiltered != settings[0].prefiltered or rs.debug != settings[0].debug for rs in settings):
            raise ValueError("all views must use the same sh_degree, scale_modifier, prefiltered and debug flags")
        mr = []
        for m, r in zip(means3D, rotations):
            mr += [m, r]
        out = _RasterizeViews.apply(tuple(settings), self._lanes_for(opacities.device), opacities, shs, scales, *mr)
        return list(out[:n]), list(out[n:])
