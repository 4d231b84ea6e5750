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


def shard_views_balanced(perm: Sequence[int], costs: Sequence[float], step: int, rank: int, world: int,
                         views_per_rank: int = 1) -> list[int]:
    """Like shard_views, but the step's world*views_per_rank views are dealt to the ranks by cost instead of by position:
    longest-processing-time-first onto the least loaded rank that still has a free slot. A synchronous data-parallel
    step lasts as long as its slowest rank, and the cost of a view (its number of (Gaussian, tile) instances,
    `num_rendered`) varies by +-10 % between the views of a scene; with 8 ranks the slowest one decides. Every rank
    computes the same assignment from the same costs (costs[v] for view index v); ties are broken by view order."""
    n = len(perm)
    base = step * world * views_per_rank
    step_views = [perm[(base + j) % n] for j in range(world * views_per_rank)]
    order = sorted(range(len(step_views)), key=lambda j: (-float(costs[step_views[j]]), j))
    load = [0.0] * world
    slots = [[] for _ in range(world)]
    for j in order:
        r = min((q for q in range(world) if len(slots[q]) < views_per_rank), key=lambda q: (load[q], q))
        slots[r].append(step_views[j])
        load[r] += float(costs[step_views[j]])
    return slots[rank]


def shard_frames(n_frames: int, rank: int, world: int) -> list[int]:
    """Frames rendered by `rank` in the forward-only all-frames render: i % world == rank, no communication."""
    return list(range(rank, n_frames, world))


_ALIGN_FLOATS = 16  # every window starts on a 64-byte boundary (the kernels use 128-bit accesses on some of them)


def _pad(n: int) -> int:
    return (n + _ALIGN_FLOATS - 1) // _ALIGN_FLOATS * _ALIGN_FLOATS


class GradBucket:
    """One flat fp32 buffer for all Gaussian-parameter gradients of a step; `views()` are the per-parameter windows the
    backward accumulates into (each starts on a 64-byte boundary, whatever the number of Gaussians), `all_reduce()`
    sums the whole buffer across ranks in one collective."""

    def __init__(self, shapes: dict[str, Sequence[int]], device, dtype=torch.float32):
        self.shapes = {k: tuple(v) for k, v in shapes.items()}
        self.numels = {k: int(torch.Size(v).numel()) for k, v in self.shapes.items()}
        total = sum(_pad(n) for n in self.numels.values())
        self.flat = torch.zeros(total, device=device, dtype=dtype)
        self._views = {}
        off = 0
        for k, n in self.numels.items():
            self._views[k] = self.flat[off:off + n].view(self.shapes[k])
            off += _pad(n)

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


def raw_param_widths(sh_degree: int = 3, trim_sh: bool = True) -> dict[str, int]:
    """Floats per Gaussian of the gradients of DAS3R's raw parameters (scene/gaussian_model.py:47-59) plus the static
    confidence: `_xyz` 3, `_rotation` 4, `_opacity` 1, `_scaling` 3, `_features_dc` 3, `_features_rest` 45, conf 1 = 60.
    With trim_sh only the SH bands of the active degree are kept: bands above it have an identically zero gradient
    (backward.cu:62-137), which is 180 of 240 bytes at degree 0 - the regime of iterations 1-2999 (train_gui.py:541-543)."""
    K = (sh_degree + 1) ** 2 if trim_sh else 16
    return {"xyz": 3, "rotation": 4, "opacity": 1, "scaling": 3, "features_dc": 3, "features_rest": 3 * (K - 1), "conf": 1}


class SlicedGradBucket:
    """Flat fp32 gradient buffer in SLICE-MAJOR order: slice c holds rows [first_c, first_c + count_c) of every
    parameter, window after window, so that a slice is one contiguous range. The backward fills slice after slice
    (das3r_raster_backward_finish_ex / das3r_frontend_backward_ex with a Gaussian range and slice-local outputs) and
    each finished slice is handed to its own all-reduce while the next one is computed: the collective overlaps the
    step's own gradient pass instead of the next step's forward, so nothing is stale. `tail` holds the small dense
    gradients (poses, ...) and travels with the last slice. Slice boundaries are multiples of 256 rows."""

    def __init__(self, P: int, widths: dict[str, int], device, n_slices: int = 1, tail: dict[str, Sequence[int]] | None = None,
                 dtype=torch.float32, allocate=None):
        self.P = int(P)
        self.widths = {k: int(w) for k, w in widths.items() if int(w) > 0}
        n_slices = max(1, min(int(n_slices), max(1, (self.P + 255) // 256)))
        per = ((self.P + n_slices - 1) // n_slices + 255) // 256 * 256
        self.slices = []
        r = 0
        while r < self.P:
            c = min(per, self.P - r)
            self.slices.append((r, c))
            r += c
        if not self.slices:
            self.slices = [(0, 0)]
        self.tail_shapes = {k: tuple(v) for k, v in (tail or {}).items()}
        total, self._win, self._bounds = 0, [], []
        for (r0, cnt) in self.slices:
            start = total
            w = {}
            for k, wd in self.widths.items():
                w[k] = (total, cnt * wd)
                total += _pad(cnt * wd)
            self._win.append(w)
            self._bounds.append([start, total])
        self._tail = {}
        for k, shp in self.tail_shapes.items():
            n = int(torch.Size(shp).numel())
            self._tail[k] = (total, n)
            total += _pad(n)
        self._bounds[-1][1] = total  # the tail rides with the last slice
        # allocate(numel) -> flat fp32 tensor: lets the caller place the bucket in symmetric memory (NvlsReducer)
        self.flat = allocate(total) if allocate is not None else torch.zeros(total, device=device, dtype=dtype)
        assert self.flat.numel() >= total and self.flat.dtype == dtype

    def slice_bounds(self, c: int) -> tuple[int, int]:
        """(offset, length) of slice c in the flat buffer, in elements; both are multiples of 16."""
        a, b = self._bounds[c]
        return a, b - a

    def window(self, c: int, name: str) -> torch.Tensor:
        off, n = self._win[c][name]
        return self.flat[off:off + n].view(self.slices[c][1], self.widths[name])

    def windows(self, c: int) -> dict[str, torch.Tensor]:
        return {k: self.window(c, k) for k in self.widths}

    def tail(self, name: str) -> torch.Tensor:
        off, n = self._tail[name]
        return self.flat[off:off + n].view(self.tail_shapes[name])

    def slice_flat(self, c: int) -> torch.Tensor:
        a, b = self._bounds[c]
        return self.flat[a:b]

    def gather(self, name: str) -> torch.Tensor:
        """[P, width] copy of one parameter's gradient (tests, optimizers that want whole tensors)."""
        return torch.cat([self.window(c, name) for c in range(len(self.slices))], dim=0)

    def all_reduce_slice(self, c: int, group=None, async_op: bool = True):
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None
        return dist.all_reduce(self.slice_flat(c), op=dist.ReduceOp.SUM, group=group, async_op=async_op)

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


class NvlsReducer:
    """Sum of a symmetric fp32 buffer over the ranks of one NVSwitch box with this library's own kernel
    (das3r_nvls_allreduce_sum: multimem.ld_reduce + multimem.st, include/das3r_comm.h) instead of NCCL. The buffer lives
    in torch's symmetric memory, which also provides the multicast mapping and the cross-rank barrier; the reductions run
    on a high-priority side stream so that a finished slice of the gradient bucket is summed while the next slice is
    still being computed.

        red = NvlsReducer(device, group)            # after init_process_group
        bucket = SlicedGradBucket(..., allocate=red.allocate)
        ... compute slice c on the current stream ...
        red.all_reduce(bucket.flat, *bucket.slice_bounds(c))   # asynchronous
        red.wait()                                   # current stream waits for every reduction issued so far

    `available` is False when the box has no multicast support (then use NCCL: SlicedGradBucket.all_reduce_slice)."""

    def __init__(self, device, group=None, n_ctas: int = 0):
        import torch.distributed._symmetric_memory as symm
        from . import load_native

        self._symm = symm
        self._C = load_native()
        self.device = torch.device(device)
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        self.n_ctas = int(n_ctas)
        self._handles = {}  # data_ptr of an allocation -> (flat tensor, symmetric-memory handle)
        self.available = False
        self.stream = torch.cuda.Stream(device=self.device, priority=-1)
        self._pending = []
        self.trace = None  # developer aid: a list collects (slice ready on the main stream, reduction done) event pairs

    def allocate(self, numel: int) -> torch.Tensor:
        flat = self._symm.empty(numel, dtype=torch.float32, device=self.device)
        flat.zero_()
        handle = self._symm.rendezvous(flat, self.group)
        self.available = int(handle.multicast_ptr) != 0
        self._handles[flat.data_ptr()] = (flat, handle)
        return flat

    def all_reduce(self, flat: torch.Tensor, offset: int, length: int) -> None:
        """Sum flat[offset:offset+length] (a buffer from `allocate`, offset and length multiples of 4) over the ranks,
        asynchronously on the reducer's stream, ordered after everything queued on the current stream so far."""
        assert self.available, "no NVSwitch multicast mapping: use NCCL"
        handle = self._handles[flat.data_ptr()][1]
        main = torch.cuda.current_stream(self.device)
        ready = torch.cuda.Event()
        ready.record(main)
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(ready)
            handle.barrier(channel=0)  # every rank has finished writing its copy of the slice
            self._C.nvls_allreduce_sum(int(handle.multicast_ptr) + int(handle.offset), int(offset), int(length), self.rank,
                                       self.world, self.n_ctas, self.device.index)
            handle.barrier(channel=0)  # every rank's share of the sums has landed in every copy
            done = torch.cuda.Event(enable_timing=self.trace is not None)
            done.record(self.stream)
        if self.trace is not None:
            ready_t = torch.cuda.Event(enable_timing=True)
            ready_t.record(main)
            self.trace.append((ready_t, done))
        self._pending.append(done)

    def wait(self) -> None:
        main = torch.cuda.current_stream(self.device)
        for ev in self._pending:
            main.wait_event(ev)
        self._pending.clear()
