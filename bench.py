#!/usr/bin/env python
"""bench.py — the headline benchmark of the rasterizer hot path (BASELINE.json metric:
"raster fwd+bwd views/s @1M Gaussians 512x288").

    python bench.py --gpus N --steps K --warmup W [--impl ours|reference|oracle-cpu]

For N > 1 the driver launches it under torchrun (one rank per GPU, NCCL); RANK/LOCAL_RANK/WORLD_SIZE come from the env.

A STEP is one pass of the hot path over one batch: every rank rasterizes `--views-per-step` disjoint camera views of
the (replicated) synthetic scene, forward + backward with a dense dL/d(colour), summing the Gaussian-parameter
gradients of its views in place into one flat fp32 bucket; for N > 1 the bucket is then summed across ranks with ONE
NCCL all-reduce per step (double-buffered, overlapping the next step). `value` = views of all ranks / device time, with
all inputs resident in HBM; `e2e` = the same metric through the public drop-in API (GaussianRasterizer + autograd) with
the per-view host inputs (camera bundle, dL/d(colour)) copied from pinned memory and the rendered image + loss read
back inside the timed region.

`--impl reference` times the reference CUDA rasterizer (oracle/_ref, built from /root/reference by
oracle/build_ref.sh) through its own pybind API in the same step structure (its fresh zero-filled gradients are added
to the bucket with torch ops, the bucket is all-reduced the same way). `--impl oracle-cpu` times the C port of the
reference algorithm on the host cores (rank 0 only). Only this file's cpu_baseline / reference legs touch oracle/.
"""
from __future__ import annotations

import argparse
from concurrent.futures import ThreadPoolExecutor
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

METRIC = "raster fwd+bwd views/s @1M Gaussians 512x288"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)  # ~1 s timed region: enough clock samples
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "oracle-cpu"])
    ap.add_argument("--views-per-step", type=int, default=8)
    ap.add_argument("--P", type=int, default=1_000_000)
    ap.add_argument("--V", type=int, default=50)
    ap.add_argument("--W", type=int, default=512)
    ap.add_argument("--H", type=int, default=288)
    ap.add_argument("--sh-degree", type=int, default=3)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-views", type=int, default=4, help="views timed for the CPU baseline sample")
    ap.add_argument("--eager-allreduce", action="store_true",
                    help="start a step's gradient all-reduce right after its backward instead of after the next "
                         "step's preprocess pass")
    ap.add_argument("--per-view-backward", action="store_true",
                    help="sum the views with the per-view accumulating backward instead of the fused multi-view pass")
    ap.add_argument("--streams", type=int, default=3,
                    help="CUDA streams the views of a step are spread over (the reference arm always uses 1: its "
                         "kernels are hard-wired to the legacy default stream)")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 100 ms while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines: list[str] = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i",
                 str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def hbm_peak() -> tuple[float, str]:
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def stage_bytes(P, R, N, T, s, fused_views=0):
    """Algorithmic bytes per launch of each stage (SURVEY.md §8d / DESIGN.md §4): compulsory traffic, each array counted
    once per pass that semantically needs it. s = 12*(D+1)^2 bytes of active SH per Gaussian. With the fused multi-view
    backward one preprocess_bwd launch serves `fused_views` views: per view it must read radii 4, the accumulator row
    48, mean 12, scale 12, rotation 16, clamp bits 1 (= 93 B per Gaussian); once per launch the SH rows (s) and the
    gradient rows it writes (opacity 4, mean 12, scale 12, rotation 16, SH s)."""
    if fused_views:
        pre_bwd = P * (93 * fused_views + 2 * s + 44)
    else:
        pre_bwd = P * (199 + 2 * s)
    return {
        "preprocess_fwd": P * (119 + s),
        "tile_scan": 8 * T,
        "scatter": 20 * P + 12 * R,
        "tile_sort": 24 * R,
        "blend_fwd": 48 * R + 20 * N,
        "blend_bwd": 40 * R + 20 * N,
        "preprocess_bwd": pre_bwd,
    }


def view_bytes(P, R, N, T, s):
    return P * (318 + 3 * s) + 124 * R + 40 * N + 8 * T


_JSON_FD = None


def emit(obj):
    """The one JSON line goes to the process's ORIGINAL stdout; everything else written to fd 1 (NCCL prints its version
    banner there from native code, whatever NCCL_DEBUG_FILE says) has been pointed at stderr by main()."""
    line = (json.dumps(obj) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(line.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, line)


def main():
    global _JSON_FD
    args = parse_args()
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    vps = args.views_per_step

    from das3r_b200.synth import synth_scene, view_inputs, grad_seed
    from das3r_b200 import sharding

    if args.impl == "oracle-cpu":
        if rank != 0:
            return
        run_cpu_arm(args)
        return

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU path)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # keep stdout for the one JSON line
        dist.init_process_group("nccl", device_id=dev)

    if args.impl == "ours":
        from das3r_b200 import load_native
        C = load_native()
    else:
        import oracle
        if not oracle.reference_available():
            if rank == 0:
                emit(({"impl": "reference", "unavailable": "oracle/_ref not built (run oracle/build_ref.sh "
                                  "where /root/reference exists)"}))
            return
        C, _ = oracle.load_reference()

    # ---- the workload: one replicated scene, per-rank view pools ------------------------------------------------
    sc = synth_scene(P=args.P, V=args.V, W=args.W, H=args.H, seed=args.seed, sh_degree=args.sh_degree)
    P, W, H, D = args.P, args.W, args.H, args.sh_degree
    N, T = W * H, ((W + 15) // 16) * ((H + 15) // 16)
    perm = torch.randperm(sc.V, generator=torch.Generator().manual_seed(1234)).tolist()
    my_views = sharding.shard_views(perm, 0, rank, world, vps)
    empty = torch.Tensor([])
    shared = None
    pool = []
    for v in my_views:
        vi = view_inputs(sc, v)
        if shared is None:
            shared = dict(shs=vi["shs"].to(dev), opac=vi["opacities"].to(dev), scales=vi["scales"].to(dev),
                          bg=vi["bg"].to(dev), view=vi["viewmatrix"].to(dev).contiguous(),
                          proj=vi["projmatrix"].to(dev).contiguous(), campos=vi["campos"].to(dev).contiguous(),
                          tanx=vi["tanfovx"], tany=vi["tanfovy"])
        pool.append(dict(means=vi["means3D"].contiguous().to(dev), rots=vi["rotations"].contiguous().to(dev)))
    dL = grad_seed(sc).to(dev)
    S = shared

    shapes = sharding.gaussian_grad_shapes(P, 16)
    # The views of a step are independent, so they are spread over S streams: the HBM-bound stages of one view
    # (preprocess, binning, gather) overlap with the issue-bound blend kernels of another. Every stream sums into its
    # own bucket (a shared one would be a read-modify-write race); the step ends by folding them into bucket 0.
    S_ = 1 if args.impl != "ours" else max(1, min(args.streams, vps))
    nbuf = 2 if world > 1 else 1
    buckets = [[sharding.GradBucket(shapes, dev) for _ in range(S_)] for _ in range(nbuf)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(S_)] if S_ > 1 else [None]
    handles = [None] * nbuf
    last_R = [0]
    workers = ThreadPoolExecutor(max_workers=S_) if S_ > 1 else None

    def fwd(pv):
        return C.rasterize_gaussians(S["bg"], pv["means"], empty, S["opac"], S["scales"], pv["rots"], 1.0, empty,
                                     S["view"], S["proj"], S["tanx"], S["tany"], H, W, S["shs"], D, S["campos"], False,
                                     False)

    def one_view(pv, bk):
        """fwd + bwd of one view with the per-view API (reference arm; ours with --per-view-backward)."""
        g = bk.views()
        R, color, radii, geom, binning, img = fwd(pv)
        last_R[0] = R
        bargs = (S["bg"], pv["means"], radii, empty, S["scales"], pv["rots"], 1.0, empty, S["view"], S["proj"],
                 S["tanx"], S["tany"], dL, S["shs"], D, S["campos"], geom, R, binning, img, False)
        if args.impl == "ours":
            outs = [empty, empty, g["opacities"], g["means3D"], empty, g["shs"], g["scales"], g["rotations"]]
            C.rasterize_gaussians_backward_into(*bargs, outs, True)
        else:
            (_m2d, _dc, d_op, d_m3, _dcov, d_sh, d_sc, d_rot) = C.rasterize_gaussians_backward(*bargs)
            g["opacities"].add_(d_op)
            g["means3D"].add_(d_m3)
            g["shs"].add_(d_sh)
            g["scales"].add_(d_sc)
            g["rotations"].add_(d_rot)

    def preprocess_all():
        """first half of the forward for all views of the step in one call (one kernel reads the shared SH rows once)"""
        return C.rasterize_gaussians_forward_preprocess(
            [(S["bg"], pv["means"], pv["rots"], S["view"], S["proj"], float(S["tanx"]), float(S["tany"]), H, W,
              S["campos"]) for pv in pool], S["opac"], S["scales"], S["shs"], D, 1.0, False, False)

    def view_blend(pv, pre_v, lane_stream=None):
        """second half of the forward + first half of the backward (backward blend) of one view; returns what the fused
        second half of the backward needs."""
        color, radii, geom, img = pre_v
        if lane_stream is not None:
            for t in (color, radii, geom, img):
                t.record_stream(lane_stream)
        R, binning = C.rasterize_gaussians_forward_render(S["bg"], color, radii, geom, img, False)
        last_R[0] = R
        acc = C.rasterize_gaussians_backward_blend(S["bg"], dL, P, geom, R, binning, img, False)
        return (pv["means"], radii, S["scales"], pv["rots"], S["view"], S["proj"], float(S["tanx"]), float(S["tany"]),
                S["campos"], geom, acc, H, W, empty, empty)

    def finish(fviews, bk):
        """second half of the backward for all views of the step: one pass, gradients summed over the views."""
        g = bk.views()
        C.rasterize_gaussians_backward_finish(fviews, S["shs"], D, 1.0,
                                              [g["opacities"], g["means3D"], g["shs"], g["scales"], g["rotations"]],
                                              False, False)

    fused_bwd = args.impl == "ours" and not args.per_view_backward
    single_stream = [False]  # the per-stage profiling pass runs the views back to back on one stream

    def step(i):
        b = i % nbuf
        if handles[b] is not None:
            handles[b].wait()
            handles[b] = None
        if S_ == 1 or single_stream[0]:
            if fused_bwd:
                pre = preprocess_all()
                finish([view_blend(pv, pre[j]) for j, pv in enumerate(pool)], buckets[b][0])
            else:
                buckets[b][0].zero_()
                for pv in pool:
                    one_view(pv, buckets[b][0])
        else:
            # one host thread per stream: the forward blocks its thread on the 16-byte read-back of num_rendered, the
            # other threads keep their streams fed meanwhile
            main = torch.cuda.current_stream()
            pre = preprocess_all() if fused_bwd else None
            launch_pending_reduce()  # ordered after this step's preprocess pass: it then runs under the blend kernels
            ev = torch.cuda.Event()
            ev.record(main)
            fviews = [None] * len(pool)

            def work(si):
                torch.cuda.set_device(dev)
                st = streams[si]
                with torch.cuda.stream(st):
                    st.wait_event(ev)
                    if not fused_bwd:
                        buckets[b][si].zero_()
                    for j in range(si, len(pool), S_):
                        if fused_bwd:
                            fviews[j] = view_blend(pool[j], pre[j], st)
                        else:
                            one_view(pool[j], buckets[b][si])

            for f in [workers.submit(work, si) for si in range(S_)]:
                f.result()
            for si, st in enumerate(streams):
                main.wait_stream(st)
            if fused_bwd:
                finish(fviews, buckets[b][0])
                for fv in fviews:  # the per-view tensors were allocated on side streams and are last read on `main`
                    for t in fv:
                        if isinstance(t, torch.Tensor) and t.is_cuda:
                            t.record_stream(main)
            else:
                for si in range(1, S_):
                    buckets[b][0].flat.add_(buckets[b][si].flat)
        if world > 1:
            # The all-reduce of this step's bucket is not started here but after the NEXT step's preprocess pass (or in
            # drain()): started right away it shares HBM with that bandwidth-bound pass; a little later it overlaps
            # the issue-bound blend kernels, which leave the memory system idle.
            launch_pending_reduce()
            pending[0] = b
            if args.eager_allreduce:
                launch_pending_reduce()

    pending = [None]

    def launch_pending_reduce():
        if world > 1 and pending[0] is not None:
            pb = pending[0]
            pending[0] = None
            handles[pb] = buckets[pb][0].all_reduce(async_op=True)

    def drain():
        launch_pending_reduce()
        for b in range(len(handles)):
            if handles[b] is not None:
                handles[b].wait()
                handles[b] = None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up, then the timed region -------------------------------------------------------------------------
    if world > 1:  # NCCL sets up its channels and buffers lazily: not inside the W warm-up steps, let alone the timed ones
        for _ in range(3):
            for bb in range(nbuf):
                buckets[bb][0].all_reduce(async_op=True).wait()
        torch.cuda.synchronize()
    for i in range(args.warmup):
        step(i)
    drain()
    barrier()
    launches0 = C.launch_count() if args.impl == "ours" else 0
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(args.steps):
        step(args.warmup + i)
    drain()
    ev1.record()
    torch.cuda.synchronize()
    elapsed_ms = ev0.elapsed_time(ev1)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = (C.launch_count() - launches0) if args.impl == "ours" else None
    elapsed_ms = sharding.all_reduce_max(elapsed_ms, dev)
    views_total = world * vps * args.steps
    value = views_total / (elapsed_ms / 1e3)
    R = int(last_R[0])
    s_bytes = 12 * (D + 1) ** 2

    out = {
        "metric": METRIC, "value": round(value, 2), "unit": "views/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(elapsed_ms / args.steps, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {
            "workload": f"C2-shaped synthetic scene: {P} Gaussians from {args.V} pointmap views, {W}x{H}, SH degree {D}, "
                        f"shs+scales/rotations inputs; step = {vps} views/rank (fwd+bwd, dense dL), gradients summed "
                        f"{'by the fused multi-view backward pass' if fused_bwd else 'in place (per-view accumulate)'} into a "
                        f"{buckets[0][0].nbytes() // (1 << 20)} MiB fp32 bucket, views spread over {S_} CUDA stream(s)"
                        + (", one NCCL all-reduce of the bucket per step (overlapped)" if world > 1 else ""),
            "views_per_step_per_rank": vps, "streams": S_, "P": P, "W": W, "H": H, "sh_degree": D, "num_rendered": R,
            "l2_policy": "inputs larger than L2 (per step each rank streams ~%d MB of distinct inputs; L2 is 126 MB)" %
                         ((P * (28 * vps + 208)) // 1_000_000),
            "parallelism": f"views sharded over {world} rank(s), Gaussians replicated",
        },
    }
    if args.impl != "ours":
        out["impl"] = "reference"
        out["cpu_baseline"] = {"value": out["value"], "unit": "views/s", "cores": 0, "kind": "reference",
                               "sample": "reference CUDA rasterizer (oracle/_ref, sm_100 build of the vendored "
                                         "diff-gaussian-rasterization with DAS3R's 0.001 near plane) on the GPU: the "
                                         "reference has no CPU implementation of this path; the CPU port is timed by "
                                         "`--impl oracle-cpu` and in the `ours` line's cpu_baseline"}
        out["e2e"] = {"value": out["value"], "unit": "views/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
        out["clocks"] = clocks
        if rank == 0:
            emit((out))
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- per-stage device times (CUDA events on the launching stream), profiled pass right after the timed region ---
    C.prof_collect(True)
    C.prof_enable(True)
    single_stream[0] = True
    prof_steps = max(2, min(args.steps, 10))
    for i in range(prof_steps):
        step(i)
    drain()
    torch.cuda.synchronize()
    single_stream[0] = False
    collected = [(name, ms, calls) for name, ms, calls in C.prof_collect(True) if calls > 0]
    stages = {name: (ms / calls) for name, ms, calls in collected}
    stage_total = {name: ms for name, ms, calls in collected}
    C.prof_enable(False)
    peak, peak_src = hbm_peak()
    sb = stage_bytes(P, R, N, T, s_bytes, fused_views=vps if fused_bwd else 0)
    kern = {k: v for k, v in stages.items() if k in sb}
    dom = max(kern, key=lambda k: stage_total[k])  # the kernel with the largest share of the step
    achieved = sb[dom] / (kern[dom] * 1e-3) / 1e9
    vb = view_bytes(P, R, N, T, s_bytes)
    per_view_ms = elapsed_ms / (vps * args.steps)
    traffic = None
    try:  # DRAM bytes per launch of that kernel from the committed `ncu --set full` capture (not measured by this run)
        traffic = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json"))).get(dom)
    except Exception:
        pass
    out["roofline"] = {
        "bound": "hbm", "kernel": dom, "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
        "frac": round(achieved / peak, 4), "traffic": traffic, "peak_source": peak_src,
        "algorithmic_bytes_per_launch": sb[dom], "avg_launch_ms": round(kern[dom], 5),
        "timing": f"CUDA events around every stage on the launching stream, {prof_steps} single-stream steps right "
                  "after the timed region",
        "stage_ms": {k: round(v, 5) for k, v in stages.items()},
        "stage_share_of_step": {k: round(v / sum(stage_total.values()), 4) for k, v in stage_total.items()},
        "sort_gkeys_per_s": round(R / (stages["tile_sort"] * 1e-3) / 1e9, 2) if "tile_sort" in stages else None,
        "stage_frac_of_hbm_peak": {k: round(sb[k] / (v * 1e-3) / 1e9 / peak, 4) for k, v in kern.items()},
        "view": {"algorithmic_bytes": vb, "achieved_gbs": round(vb / (per_view_ms * 1e-3) / 1e9, 1),
                 "frac": round(vb / (per_view_ms * 1e-3) / 1e9 / peak, 4),
                 "roofline_views_per_s_per_gpu": round(peak * 1e9 / vb, 1)},
    }
    out["gpu_launches"] = int(launches)
    out["clocks"] = clocks

    # ---- e2e: public API (GaussianRasterizer + autograd), per-view host inputs from pinned memory -------------------
    if not args.no_e2e:
        out["e2e"] = run_e2e(args, sc, dev, world, rank, pool, S, dL, my_views)
    # ---- CPU baseline: the C port of the reference algorithm on the host cores (rank 0, N = 1 only) -----------------
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(args, sc)
    if rank == 0:
        emit((out))
    if world > 1:
        dist.destroy_process_group()


def run_e2e(args, sc, dev, world, rank, pool, S, dL, my_views):
    """views/s through the public Python API with HOST inputs. Per view, inside the timed region: H2D (pinned) of the
    camera bundle (viewmatrix 16 + projmatrix 16 + campos 3 + bg 3 floats) and of dL/d(colour) [3,H,W]; forward; loss;
    autograd backward; D2H of the rendered image; D2H of the step's loss. The Gaussian parameters are the resident
    model state (like weights). For N > 1 the leaf gradients are all-reduced once per step.

    Two surfaces are timed: `MultiViewRasterizer` (das3r_b200/multiview.py: all views of the step in one autograd node -
    the headline `value`) and the drop-in per-view `GaussianRasterizer` (`per_view_api`)."""
    from das3r_b200.rasterizer import GaussianRasterizationSettings, GaussianRasterizer
    from das3r_b200.multiview import MultiViewRasterizer
    from das3r_b200 import sharding

    W, H, D, vps = args.W, args.H, args.sh_degree, args.views_per_step
    cam_host = torch.cat([S["view"].flatten().cpu(), S["proj"].flatten().cpu(), S["campos"].cpu(), S["bg"].cpu()]).pin_memory()
    dL_host = dL.cpu().pin_memory()
    cam_dev = [torch.empty_like(cam_host, device=dev) for _ in range(vps)]
    dL_dev = [torch.empty_like(dL_host, device=dev) for _ in range(vps)]
    img_host = [torch.empty(3, H, W).pin_memory() for _ in range(vps)]
    loss_host = torch.empty(1).pin_memory()
    leaves = dict(shs=S["shs"].clone().requires_grad_(True), opac=S["opac"].clone().requires_grad_(True),
                  scales=S["scales"].clone().requires_grad_(True))
    per_view = [dict(means=pv["means"].clone().requires_grad_(True), rots=pv["rots"].clone().requires_grad_(True))
                for pv in pool]
    means2D = torch.zeros(args.P, 3, device=dev, requires_grad=True)
    all_leaves = list(leaves.values()) + [means2D] + [x for pv in per_view for x in pv.values()]
    mv = MultiViewRasterizer(streams=max(1, min(args.streams, vps)))
    copy_stream = torch.cuda.Stream(device=dev)

    def settings_for(v):
        c = cam_dev[v]
        return GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=S["tanx"], tanfovy=S["tany"], bg=c[35:38], scale_modifier=1.0,
            viewmatrix=c[0:16].view(4, 4), projmatrix=c[16:32].view(4, 4), sh_degree=D, campos=c[32:35],
            prefiltered=False, debug=False)

    def reduce_grads():
        if world > 1:
            for t in (leaves["shs"], leaves["opac"], leaves["scales"], per_view[0]["means"], per_view[0]["rots"]):
                dist.all_reduce(t.grad)

    def step_multi():
        main = torch.cuda.current_stream()
        for t in all_leaves:
            t.grad = None  # optimizer.zero_grad(set_to_none=True)
        for v in range(vps):
            cam_dev[v].copy_(cam_host, non_blocking=True)
        copy_stream.wait_stream(main)  # the previous step is done with dL_dev
        with torch.cuda.stream(copy_stream):
            for v in range(vps):
                dL_dev[v].copy_(dL_host, non_blocking=True)  # only needed by the loss: overlaps the forwards
        colors, _radii = mv([settings_for(v) for v in range(vps)], [pv["means"] for pv in per_view],
                            [pv["rots"] for pv in per_view], leaves["opac"], leaves["shs"], leaves["scales"])
        main.wait_stream(copy_stream)
        loss = sum((c * dL_dev[v]).sum() for v, c in enumerate(colors))
        copy_stream.wait_stream(main)
        with torch.cuda.stream(copy_stream):  # images go home while the backward runs
            for v, c in enumerate(colors):
                img_host[v].copy_(c.detach(), non_blocking=True)
                c.record_stream(copy_stream)
        loss.backward()
        loss_host.copy_(loss.detach().view(1), non_blocking=True)
        reduce_grads()
        main.wait_stream(copy_stream)
        main.synchronize()  # the step's images and loss are on the host now

    def step_per_view():
        for t in all_leaves:
            if t.grad is not None:
                t.grad.zero_()
        total = None
        for v, pv in enumerate(per_view):
            cam_dev[v].copy_(cam_host, non_blocking=True)
            dL_dev[v].copy_(dL_host, non_blocking=True)
            color, _radii = GaussianRasterizer(settings_for(v))(
                means3D=pv["means"], means2D=means2D, opacities=leaves["opac"], shs=leaves["shs"],
                scales=leaves["scales"], rotations=pv["rots"])
            loss = (color * dL_dev[v]).sum()
            loss.backward()
            img_host[v].copy_(color.detach(), non_blocking=True)
            total = loss.detach() if total is None else total + loss.detach()
        loss_host.copy_(total.view(1), non_blocking=True)
        reduce_grads()
        torch.cuda.current_stream().synchronize()

    def timed(step, steps):
        for _ in range(2):
            step()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = sharding.all_reduce_max(e0.elapsed_time(e1), dev)
        if world > 1:
            dist.barrier()
        return round(world * vps * steps / (ms / 1e3), 2)

    # The same step started one stage earlier: raw model parameters + camera poses -> fused front-end
    # (das3r_b200/frontend.py: what gaussian_renderer/__init__.py:83-126 does with ~10 torch kernels per view) ->
    # MultiViewRasterizer; the backward reaches the raw parameters and the poses.
    from das3r_b200.frontend import render_inputs
    raw = dict(xyz=per_view[0]["means"].detach().clone(), rotation=per_view[0]["rots"].detach().clone(),
               opacity=torch.logit(S["opac"].clamp(1e-6, 1 - 1e-6)), scaling=torch.log(S["scales"]),
               features_dc=S["shs"][:, :1].contiguous(), features_rest=S["shs"][:, 1:].contiguous(),
               conf_static=torch.ones(args.P, device=dev))
    raw = {k: v.requires_grad_(True) for k, v in raw.items()}
    poses_host = torch.tensor([[1.0, 0, 0, 0, 0, 0, 0]] * vps).pin_memory()  # identity: the scene stays in view
    poses_dev = torch.empty(vps, 7, device=dev, requires_grad=True)

    def step_frontend():
        main = torch.cuda.current_stream()
        for t in list(raw.values()) + [poses_dev]:
            t.grad = None
        with torch.no_grad():
            poses_dev.copy_(poses_host, non_blocking=True)
        for v in range(vps):
            cam_dev[v].copy_(cam_host, non_blocking=True)
        copy_stream.wait_stream(main)
        with torch.cuda.stream(copy_stream):
            for v in range(vps):
                dL_dev[v].copy_(dL_host, non_blocking=True)
        fe = render_inputs(raw["xyz"], raw["rotation"], raw["opacity"], raw["scaling"], raw["features_dc"],
                           raw["features_rest"], raw["conf_static"], None, poses_dev)
        colors, _radii = mv([settings_for(v) for v in range(vps)], fe.means3D, fe.rotations, fe.opacity, fe.shs, fe.scales)
        main.wait_stream(copy_stream)
        loss = sum((c * dL_dev[v]).sum() for v, c in enumerate(colors))
        copy_stream.wait_stream(main)
        with torch.cuda.stream(copy_stream):
            for v, c in enumerate(colors):
                img_host[v].copy_(c.detach(), non_blocking=True)
                c.record_stream(copy_stream)
        loss.backward()
        loss_host.copy_(loss.detach().view(1), non_blocking=True)
        if world > 1:
            for t in raw.values():
                dist.all_reduce(t.grad)
        main.wait_stream(copy_stream)
        main.synchronize()

    steps = max(3, min(args.steps // 3, 40))
    v_multi = timed(step_multi, steps)
    v_single = timed(step_per_view, steps)
    v_front = timed(step_frontend, steps)
    return {"value": v_multi, "unit": "views/s",
            "h2d_bytes_per_step": vps * (38 * 4 + 3 * H * W * 4), "d2h_bytes_per_step": vps * (3 * H * W * 4) + 4,
            "steps": steps,
            "what": "das3r_b200.multiview.MultiViewRasterizer (all views of the step in one autograd node: forwards and "
                    "backward blends on CUDA streams, one fused gradient pass) + autograd; per view H2D (pinned) of the "
                    "camera bundle and dL/dcolour and D2H of the rendered image, per step D2H of the loss; Gaussian "
                    "parameters resident (model state)",
            "from_raw_parameters": {"value": v_front, "unit": "views/s",
                                    "what": "same step started from the raw model parameters and camera poses: fused "
                                            "front-end (das3r_b200.frontend.render_inputs) + MultiViewRasterizer; "
                                            "gradients reach the raw parameters and the poses"},
            "per_view_api": {"value": v_single, "unit": "views/s",
                             "what": "the drop-in GaussianRasterizer.forward + autograd backward called once per view, "
                                     "single stream, same copies"}}


def cpu_baseline(args, sc):
    """C port of the reference algorithm (oracle/raster_oracle.c, OpenMP) on a bounded sample of the same workload."""
    import oracle
    from das3r_b200.synth import view_inputs, grad_seed

    oracle.build()
    dL = grad_seed(sc)
    n = max(1, args.cpu_views)
    vi = view_inputs(sc, 0, sh_degree=args.sh_degree)
    fw = oracle.forward(**vi)  # warm-up (page-in, thread pool)
    # exclude input construction: time only the oracle calls
    t = 0.0
    for v in range(n):
        vi = view_inputs(sc, v % sc.V, sh_degree=args.sh_degree)
        a = time.perf_counter()
        fw = oracle.forward(**vi)
        oracle.backward(fw, dL, **vi)
        t += time.perf_counter() - a
    return {"value": round(n / t, 4), "unit": "views/s", "cores": os.cpu_count(), "kind": "port",
            "sample": f"{n} views (fwd+bwd) of the same {args.P}-Gaussian {args.W}x{args.H} scene through the C port of "
                      f"the reference algorithm (oracle/raster_oracle.c, OpenMP on all {os.cpu_count()} host cores)"}


def run_cpu_arm(args):
    from das3r_b200.synth import synth_scene

    sc = synth_scene(P=args.P, V=args.V, W=args.W, H=args.H, seed=args.seed, sh_degree=args.sh_degree)
    cb = cpu_baseline(args, sc)
    emit(({"impl": "oracle-cpu", "metric": METRIC, "value": cb["value"], "unit": "views/s", "n_gpus": 0,
                      "steps": args.cpu_views, "warmup": 1, "ms_per_step": round(1e3 / cb["value"], 2),
                      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                      "data": "synthetic", "config": {"workload": cb["sample"]}, "cpu_baseline": cb,
                      "e2e": {"value": cb["value"], "unit": "views/s", "h2d_bytes_per_step": 0,
                              "d2h_bytes_per_step": 0}}))


if __name__ == "__main__":
    main()
