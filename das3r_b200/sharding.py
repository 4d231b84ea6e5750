"""Multi-GPU plan of the hot path: one process per GPU, every rank holds a replica of the Gaussians,
camera views are sharded, and the Gaussian-parameter gradients of a step are summed with ONE all-reduce over a flat
fp32 buffer (NCCL over NVLink on the GPU box, gloo in the CPU tests). Frame rendering (render.py, reference
render.py:74-86) shards frames with no communication.

The reference itself is single-GPU (scripts/rendering_sintel.sh:3,57); this module adds the sharding, it replaces
nothing.
"""
from __future__ import annotations

from typing import Iterable, Sequence

import torch
import torch.distributed as dist


def shard_views(perm: Sequence[int], step: int, rank: int, world: int, views_per_rank: int = 1) -> list[int]:
    """Views rank `rank` rasterizes at `step`: a disjoint slice of the step's world*views_per_rank views, taken from
    the shared permutation `perm` (wrapping around), so that all ranks together consume consecutive entries."""
    n = len(perm)
    base = step * world * views_per_rank
    return [perm[(base + rank * views_per_rank + j) % n] for j in range(views_per_rank)]


def shard_frames(n_frames: int, rank: int, world: int) -> list[int]:
    """Frames rendered by `rank` in the forward-only all-frames render: i % world == rank, no communication."""
    return list(range(rank, n_frames, world))


class GradBucket:
    """One flat fp32 buffer for all Gaussian-parameter gradients of a step; `views()` are the per-parameter windows the
    backward accumulates into, `all_reduce()` sums the whole buffer across ranks in one collective."""

    def __init__(self, shapes: dict[str, Sequence[int]], device, dtype=torch.float32):
        self.shapes = {k: tuple(v) for k, v in shapes.items()}
        self.numels = {k: int(torch.Size(v).numel()) for k, v in self.shapes.items()}
        total = sum(self.numels.values())
        self.flat = torch.zeros(total, device=device, dtype=dtype)
        self._views = {}
        off = 0
        for k, n in self.numels.items():
            self._views[k] = self.flat[off:off + n].view(self.shapes[k])
            off += n

    def views(self) -> dict[str, torch.Tensor]:
        return self._views

    def zero_(self) -> None:
        self.flat.zero_()

    def accumulate(self, grads: dict[str, torch.Tensor]) -> None:
        for k, g in grads.items():
            if g is not None and k in self._views:
                self._views[k].add_(g.view(self.shapes[k]))

    def all_reduce(self, group=None, async_op: bool = False):
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None
        return dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)

    def nbytes(self) -> int:
        return self.flat.numel() * self.flat.element_size()


def gaussian_grad_shapes(P: int, M: int = 16) -> dict[str, tuple[int, ...]]:
    """Shapes of the per-Gaussian gradients the rasterizer returns for the SH + scale/rotation input variant
    (rasterize_points.cu:151-159): 59 floats = 236 bytes per Gaussian at M = 16."""
    return {"means3D": (P, 3), "shs": (P, M, 3), "opacities": (P, 1), "scales": (P, 3), "rotations": (P, 4)}


def all_reduce_max(value: float, device) -> float:
    """Max over ranks of a scalar (used for device-side timings: a multi-GPU step takes as long as its slowest rank)."""
    t = torch.tensor([value], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
