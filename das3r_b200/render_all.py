"""render.py's all-frames render (reference render.py:74-124: `render_set` renders every camera of the trajectory with the
trained model and saves the frames) on tensors: no dataset, no ply reader, no PNG writer - the frames come back as
tensors (or go to a caller-supplied sink, e.g. pinned host buffers), `frames_per_batch` at a time through the fused
front-end and `MultiViewRasterizer` under `torch.no_grad()`; the reference's loop keeps autograd state per frame that
nothing uses (render.py:102-107 only wraps the scene loading). Frames are independent: with several ranks each takes
`sharding.shard_frames(n, rank, world)` and nothing is communicated.

After `load_ply` the model's `_conf_static` is per Gaussian ([P,1], scene/gaussian_model.py:381,:414) and `render_test`
multiplies the opacity by it directly (gaussian_renderer/__init__.py:226); that is `conf_index=None` here."""
from __future__ import annotations

from typing import Callable, Iterable, Sequence

import torch

from . import sharding
from .frontend import render_inputs
from .multiview import MultiViewRasterizer
from .train import GaussianState, ViewGeometry


@torch.no_grad()
def render_frames(state: GaussianState, geom: ViewGeometry, poses: torch.Tensor, frames_per_batch: int = 8,
                  frames: Iterable[int] | None = None, sink: Callable[[int, torch.Tensor], None] | None = None,
                  sh_degree: int | None = None) -> list[torch.Tensor]:
    """poses [F,7] (quaternion r,x,y,z + translation per frame = `get_tensor_from_camera(view.world_view_transform.T)`,
    render.py:78). Renders `frames` (default: all) and returns the images in that order, or hands each to
    `sink(frame_index, image)` and returns []. `sh_degree` defaults to the model's active degree (3 after load_ply,
    gaussian_model.py:418)."""
    if not poses.is_cuda:
        raise RuntimeError("poses must be a CUDA tensor: the renderer has no CPU path")
    idx = list(range(poses.shape[0])) if frames is None else list(frames)
    D = state.active_sh_degree if sh_degree is None else sh_degree
    mv = MultiViewRasterizer()
    out = []
    conf = state.conf_static.reshape(-1)
    b = max(1, min(int(frames_per_batch), 16))
    for i0 in range(0, len(idx), b):
        chunk = idx[i0:i0 + b]
        fe = render_inputs(state.xyz, state.rotation, state.opacity, state.scaling, state.features_dc, state.features_rest,
                           conf, state.conf_index, poses[chunk].contiguous(), sh_degree=D)
        colors, _radii = mv([geom.settings(D)] * len(chunk), fe.means3D, fe.rotations, fe.opacity, fe.shs, fe.scales)
        for f, c in zip(chunk, colors):
            if sink is not None:
                sink(f, c)
            else:
                out.append(c)
    return out


def render_my_frames(state: GaussianState, geom: ViewGeometry, poses: torch.Tensor, rank: int, world: int, **kw):
    """The frames of this rank in the frame-sharded render (i % world == rank): -> (frame indices, images)."""
    mine = sharding.shard_frames(poses.shape[0], rank, world)
    return mine, render_frames(state, geom, poses, frames=mine, **kw)
