#!/usr/bin/env python
"""bench.py — the headline benchmark of the rasterizer hot path (BASELINE.json metric:
"raster fwd+bwd views/s @1M Gaussians 512x288").

    python bench.py --gpus N --steps K --warmup W [--impl ours|reference|oracle-cpu] [--config C1..C5] [--sh-degree D]

For N > 1 the driver launches it under torchrun (one rank per GPU, NCCL); RANK/LOCAL_RANK/WORLD_SIZE come from the env.

A STEP is one data-parallel DAS3R training step without the loss and the optimizer: every rank takes
`--views-per-step` disjoint camera views of the (replicated) synthetic scene and runs

    ours       raw parameters + the views' poses -> fused front-end (what render() does before the rasterizer call)
               -> rasterizer forward of all views -> rasterizer backward with a dense dL/d(colour)
               -> front-end backward -> gradients of _xyz/_rotation/_opacity/_scaling/_features_*/conf and the poses,
               summed over the rank's views, in one flat fp32 bucket.  Every stage is ONE launch for all views.
    reference  the reference CUDA rasterizer (oracle/_ref) forward + backward per view on pre-posed inputs; its fresh
               gradient tensors are added into a bucket with torch ops (no front-end: less work per step than `ours`).

For N > 1 the bucket is summed across ranks inside the step: the gradient pass produces it slice by slice and every
finished slice starts its own NCCL all-reduce, which overlaps the REMAINING slices of the same step; the step ends when
the last all-reduce has completed (the next step's forward would need the reduced gradients - nothing is stale).

`value` = views of all ranks / device time, with all inputs resident in HBM; `e2e` = the same step through the public
Python API (das3r_b200.frontend.render_inputs + MultiViewRasterizer + autograd) with the per-step host inputs (poses,
camera bundles, dL/d(colour)) copied from pinned memory and the rendered images + loss read back inside the timed
region.

`--impl reference` times the reference CUDA rasterizer (oracle/_ref, built from /root/reference by
oracle/build_ref.sh) through its own pybind API. `--impl oracle-cpu` times the C port of the reference algorithm on the
host cores (rank 0 only). Only this file's cpu_baseline / reference legs touch oracle/.
"""
from __future__ import annotations

import argparse
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
    ap.add_argument("--config", default="C2", choices=["C1", "C2", "C3", "C4", "C5"],
                    help="BASELINE.json config shape (C2 = the metric's; C5 is forward-only like render.py)")
    ap.add_argument("--views-per-step", type=int, default=8)
    ap.add_argument("--P", type=int, default=None)
    ap.add_argument("--V", type=int, default=None)
    ap.add_argument("--W", type=int, default=None)
    ap.add_argument("--H", type=int, default=None)
    ap.add_argument("--sh-degree", type=int, default=3)
    ap.add_argument("--seed", type=int, default=None)
    ap.add_argument("--slices", type=int, default=0,
                    help="gradient bucket slices: slice c is all-reduced while slice c+1 is computed (N > 1); 0 = the "
                         "measured best: 4 with NCCL (2 and 4 ranks), 2 with the in-switch kernel (8 ranks)")
    ap.add_argument("--nccl-normal-priority", action="store_true",
                    help="NCCL's stream at normal priority (default: high, so that a slice's all-reduce is not queued "
                         "behind the blocks of the next slice's gradient kernels)")
    ap.add_argument("--allreduce", default="auto", choices=["auto", "nccl", "nvls"],
                    help="how the bucket is summed across ranks: this library's in-switch kernel (multimem.ld_reduce / "
                         "multimem.st on torch symmetric memory) or NCCL; auto = nvls when the box supports multicast")
    ap.add_argument("--nvls-ctas", type=int, default=32, help="thread blocks of the in-switch all-reduce kernel")
    ap.add_argument("--unbalanced", action="store_true",
                    help="shard the step's views by position in the permutation instead of by cost (num_rendered)")
    ap.add_argument("--sweep-slices", default="",
                    help="developer option: comma-separated slice counts timed one after the other in this process "
                         "(extra lines on stderr); the JSON line reports --slices")
    ap.add_argument("--no-trim", action="store_true", help="all-reduce all 16 SH bands even when the degree is below 3")
    ap.add_argument("--forward-only", action="store_true", help="forward only (C5 sets it)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-views", type=int, default=4, help="views timed for the CPU baseline sample")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 20 ms while the timed region runs."""

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
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20", "-i",
                 str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
            t0 = time.time()
            while not self.lines and time.time() - t0 < 2.0:  # the first sample takes a while: have it before timing
                time.sleep(0.01)
            self.lines.clear()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
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


def stage_bytes(P, Rs, N, T, s, n, V_fe, fwd_only):
    """Algorithmic bytes PER LAUNCH of each stage (SURVEY.md §8d / DESIGN.md §4): compulsory traffic, each array counted
    once per pass that semantically needs it. Every stage is one launch for the step's n views. s = 12*(D+1)^2 bytes of
    active SH per Gaussian, Rs = per-view num_rendered.
      preprocess_fwd   the shared SH rows once + per view inputs/radii/saved geometry: P*(s + 127 n)
      preprocess_bwd   per view radius 4 + accumulator row 48 + mean 12 + scale 12 + rotation 16 + clamp bits 1 = 93;
                       once the SH rows (s) and the gradient rows written (opacity 4, scale 12, SH s) + per view the
                       mean/rotation gradients (28): P*((93 + 28) n + 2 s + 16)
      frontend_fwd/bwd the front-end's own arrays, SH copy limited to the active bands"""
    R = sum(Rs)
    out = {
        "preprocess_fwd": P * (s + 127 * n),
        "tile_scan": 8 * T * n,
        "scatter": 20 * P * n + 12 * R,
        "tile_sort": 24 * R,
        "blend_fwd": 48 * R + 20 * N * n,
        "frontend_fwd": P * (56 + 2 * s + 28 * V_fe),
    }
    if not fwd_only:
        out.update({
            "blend_bwd": 40 * R + 20 * N * n,
            "preprocess_bwd": P * (121 * n + 2 * s + 16),
            "frontend_bwd": P * (72 + 48 + 2 * s + 28 * V_fe),
        })
    return out


def view_bytes(P, R, N, T, s, fwd_only=False):
    if fwd_only:
        return P * (119 + s) + 84 * R + 20 * N + 8 * T
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


def resolve_config(args):
    from das3r_b200.synth import CONFIGS
    c = dict(CONFIGS[args.config])
    for k, a in (("P", args.P), ("V", args.V), ("W", args.W), ("H", args.H), ("seed", args.seed)):
        if a is not None:
            c[k] = a
    c["sh_degree"] = args.sh_degree
    c.pop("n_poses", None)
    if args.config == "C5":
        args.forward_only = True
    return c


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

    cfg = resolve_config(args)
    args.P, args.V, args.W, args.H, args.seed = cfg["P"], cfg["V"], cfg["W"], cfg["H"], cfg["seed"]

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
        opts = None
        if not args.nccl_normal_priority:
            try:
                opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)
            except Exception:
                opts = None
        if opts is not None:
            dist.init_process_group("nccl", device_id=dev, pg_options=opts)
        else:
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
    # multi-million-point scenes: the 3-NN distances that size the Gaussians come from the GPU (each arm uses its own
    # distCUDA2 - the GPU tests pin ours to the reference's), the CPU k-d tree takes minutes there
    dist2_fn = None
    if args.P >= 3_000_000:
        if args.impl == "ours":
            dist2_fn = lambda xyz: C.distCUDA2(xyz.to(dev)).cpu()
        else:
            dist2_fn = lambda xyz: oracle.load_reference()[1].distCUDA2(xyz.to(dev)).cpu()
    sc = synth_scene(P=args.P, V=args.V, W=args.W, H=args.H, seed=args.seed, sh_degree=args.sh_degree, dist2_fn=dist2_fn)
    P, W, H, D = args.P, args.W, args.H, args.sh_degree
    K = (D + 1) ** 2
    N, T = W * H, ((W + 15) // 16) * ((H + 15) // 16)
    fwd_only = bool(args.forward_only)
    perm = torch.randperm(sc.V, generator=torch.Generator().manual_seed(1234)).tolist()
    empty = torch.Tensor([])
    vi0 = view_inputs(sc, 0)
    S = dict(shs=vi0["shs"].to(dev), opac=vi0["opacities"].to(dev), scales=vi0["scales"].to(dev), bg=vi0["bg"].to(dev),
             view=vi0["viewmatrix"].to(dev).contiguous(), proj=vi0["projmatrix"].to(dev).contiguous(),
             campos=vi0["campos"].to(dev).contiguous(), tanx=float(vi0["tanfovx"]), tany=float(vi0["tanfovy"]))
    dL = grad_seed(sc).to(dev)
    last_R = [0]
    all_R = [[0] * vps]

    # ---- which views this rank takes: the step's world*vps views dealt by cost (num_rendered of every view, computed
    # once here by each arm with its own forward - the counts are bit-identical, so both arms get the same assignment)
    step_views = sorted(set(perm[j % sc.V] for j in range(world * vps)))
    costs = [0.0] * sc.V
    if world > 1 and not args.unbalanced:
        for v in step_views:
            vi = view_inputs(sc, v)
            a_ = (S["bg"], vi["means3D"].contiguous().to(dev), empty, S["opac"], S["scales"],
                  vi["rotations"].contiguous().to(dev), 1.0, empty, S["view"], S["proj"], S["tanx"], S["tany"], H, W,
                  S["shs"], D, S["campos"], False, False)
            costs[v] = float(C.rasterize_gaussians(*a_)[0])
        my_views = sharding.shard_views_balanced(perm, costs, 0, rank, world, vps)
    else:
        my_views = sharding.shard_views(perm, 0, rank, world, vps)
    torch.cuda.empty_cache()

    if args.impl == "ours":
        # raw model parameters (what GaussianModel holds) + this rank's poses: the front-end poses the Gaussians
        raw = dict(xyz=sc.xyz.to(dev).contiguous(), rotation=sc.rot.to(dev).contiguous(),
                   opacity=torch.logit(sc.opacity.clamp(1e-6, 1 - 1e-6)).to(dev).contiguous(),
                   scaling=torch.log(sc.scales).to(dev).contiguous(),
                   features_dc=sc.shs[:, :1].to(dev).contiguous(), features_rest=sc.shs[:, 1:].to(dev).contiguous(),
                   conf_static=torch.ones(P, device=dev))
        poses = sc.poses[my_views].to(dev).contiguous()
        ci = torch.empty(0, dtype=torch.int64)
        fi = torch.empty(0, dtype=torch.int64)
        Kb = K if not args.no_trim else 16
        widths = sharding.raw_param_widths(D, trim_sh=not args.no_trim)
        want_nvls = args.allreduce == "nvls" or (args.allreduce == "auto" and world >= 8)
        n_slices = (args.slices if args.slices > 0 else (2 if want_nvls else 4)) if (world > 1 and not fwd_only) else 1
        reducer = None
        # auto: the in-switch kernel from 8 ranks on (measured on B200 HGX, 240 MB alone: 0.55 ms vs NCCL 0.65 ms at 8
        # ranks; 0.57 vs 0.62 at 4, where the step as a whole is faster with NCCL; with 2 ranks a multicast reduction moves
        # more bytes per link than NCCL's ring: 0.61 vs 0.48 ms)
        if world > 1 and not fwd_only and (args.allreduce == "nvls" or (args.allreduce == "auto" and world >= 8)):
            try:
                reducer = sharding.NvlsReducer(dev, n_ctas=args.nvls_ctas)
            except Exception as ex:  # symmetric memory unavailable in this build
                if args.allreduce == "nvls":
                    raise
                sys.stderr.write(f"[bench] symmetric memory unavailable ({ex}); using NCCL\n")
                reducer = None

        def make_bucket(ns):
            nonlocal reducer
            if reducer is not None:
                try:
                    b_ = sharding.SlicedGradBucket(P, widths, dev, ns, tail={"poses": (vps, 7)}, allocate=reducer.allocate)
                    ok = reducer.available
                except Exception as ex:  # symmetric-memory rendezvous refused on this box
                    sys.stderr.write(f"[bench] symmetric memory rendezvous failed ({ex}); using NCCL\n")
                    b_, ok = None, False
                # every rank must take the same branch: the collective below would otherwise deadlock
                flag = torch.tensor([1.0 if ok else 0.0], device=dev)
                dist.all_reduce(flag, op=dist.ReduceOp.MIN)
                if float(flag.item()) > 0.5:
                    return b_
                if args.allreduce == "nvls":
                    raise RuntimeError("no NVSwitch multicast support on this box")
                reducer = None
                if b_ is not None:
                    return b_
            return sharding.SlicedGradBucket(P, widths, dev, ns, tail={"poses": (vps, 7)})

        holder = {"bucket": make_bucket(n_slices)}
        bucket = holder["bucket"]
        o = dict(device=dev, dtype=torch.float32)
        # intermediates of the backward, allocated once: the rasterizer's gradients w.r.t. its own inputs
        g_opac, g_scales, g_shs = torch.empty(P, 1, **o), torch.empty(P, 3, **o), torch.empty(P, Kb, 3, **o)
        g_msum, g_rsum = torch.empty(P, 3, **o), torch.empty(P, 4, **o)  # unused (every view has its own)
        g_means = [torch.empty(P, 3, **o) for _ in range(vps)]
        g_rots = [torch.empty(P, 4, **o) for _ in range(vps)]
        fe_scratch = torch.empty(((P + 255) // 256) * vps * 16 * 4 + 256, dtype=torch.uint8, device=dev)
        rest_w = widths.get("features_rest", 0) // 3

        def step(_i):
            opac, scales, shs, means, rots = C.frontend_forward_ex(
                raw["xyz"], raw["rotation"], raw["opacity"], raw["scaling"], raw["features_dc"], raw["features_rest"],
                raw["conf_static"], ci, poses, D, fi)
            pre = C.rasterize_gaussians_forward_preprocess(
                [(S["bg"], means[v], rots[v], S["view"], S["proj"], S["tanx"], S["tany"], H, W, S["campos"])
                 for v in range(vps)], opac, scales, shs, D, 1.0, False, False)
            res = C.rasterize_gaussians_forward_render_batch(
                [(S["bg"], pre[v][0], pre[v][1], pre[v][2], pre[v][3]) for v in range(vps)], False)
            all_R[0] = [r[0] for r in res]
            last_R[0] = res[-1][0]
            if fwd_only:
                return [p[0] for p in pre]
            accs = C.rasterize_gaussians_backward_blend_batch(
                [(S["bg"], dL, pre[v][2], res[v][0], res[v][1], pre[v][3]) for v in range(vps)], P, False)
            fviews = [(means[v], pre[v][1], scales, rots[v], S["view"], S["proj"], S["tanx"], S["tany"], S["campos"],
                       pre[v][2], accs[v], H, W, g_means[v], g_rots[v]) for v in range(vps)]
            handles = []
            bucket = holder["bucket"]
            last = len(bucket.slices) - 1
            for c, (first, count) in enumerate(bucket.slices):
                C.rasterize_gaussians_backward_finish_ex(fviews, shs, D, 1.0, [g_opac, g_msum, g_shs, g_scales, g_rsum],
                                                         False, False, Kb, first, count)
                w = bucket.windows(c)
                rest = w["features_rest"].view(count, rest_w, 3) if rest_w else torch.empty(0, **o)
                C.frontend_backward_ex(raw["xyz"], raw["rotation"], raw["opacity"], raw["scaling"], raw["conf_static"],
                                       ci, poses, g_opac, g_scales, g_shs, g_means, g_rots, D, fi,
                                       [w["xyz"], w["rotation"], w["opacity"], w["scaling"],
                                        w["features_dc"].view(count, 1, 3), rest, w["conf"].view(count),
                                        bucket.tail("poses")], fe_scratch, first, count, c == last)
                if world > 1 and not holder.get("no_allreduce"):
                    if reducer is not None:
                        reducer.all_reduce(bucket.flat, *bucket.slice_bounds(c))
                    else:
                        handles.append(bucket.all_reduce_slice(c, async_op=True))
            # the step is over when its gradients are reduced: nothing is carried into the next step
            if reducer is not None:
                reducer.wait()
            for h in handles:
                h.wait()
            return None

        bucket_bytes = bucket.nbytes()
        impl_detail = {
            "what": "fused front-end -> rasterizer (every stage one launch for all views of the step) -> front-end "
                    "backward into a slice-major fp32 bucket of raw-parameter gradients",
            "bucket_slices": len(bucket.slices), "bucket_bytes": bucket_bytes,
            "sh_bands_in_bucket": Kb, "streams": 1, "host_threads": 1,
            "allreduce": ("one all-reduce per bucket slice, started when the slice is complete, all waited on before the "
                          "step ends; " + ("das3r_nvls_allreduce_sum (multimem.ld_reduce + multimem.st on torch symmetric "
                                           "memory)" if reducer is not None else "NCCL")) if world > 1 else "none (1 rank)",
        }
    else:
        pool = []
        for v in my_views:
            vi = view_inputs(sc, v)
            pool.append(dict(means=vi["means3D"].contiguous().to(dev), rots=vi["rotations"].contiguous().to(dev)))
        shapes = sharding.gaussian_grad_shapes(P, 16)
        bucket = sharding.GradBucket(shapes, dev)

        def step(_i):
            g = bucket.views()
            if not fwd_only:
                bucket.zero_()
            for pv in pool:
                R, color, radii, geom, binning, img = C.rasterize_gaussians(
                    S["bg"], pv["means"], empty, S["opac"], S["scales"], pv["rots"], 1.0, empty, S["view"], S["proj"],
                    S["tanx"], S["tany"], H, W, S["shs"], D, S["campos"], False, False)
                last_R[0] = R
                if fwd_only:
                    continue
                (_m2d, _dc, d_op, d_m3, _dcov, d_sh, d_sc, d_rot) = C.rasterize_gaussians_backward(
                    S["bg"], pv["means"], radii, empty, S["scales"], pv["rots"], 1.0, empty, S["view"], S["proj"],
                    S["tanx"], S["tany"], dL, S["shs"], D, S["campos"], geom, R, binning, img, False)
                g["opacities"].add_(d_op)
                g["means3D"].add_(d_m3)
                g["shs"].add_(d_sh)
                g["scales"].add_(d_sc)
                g["rotations"].add_(d_rot)
            if world > 1 and not fwd_only:
                bucket.all_reduce(async_op=False)
            return None

        bucket_bytes = bucket.nbytes()
        impl_detail = {
            "what": "reference CUDA rasterizer (oracle/_ref) forward + backward per view on pre-posed inputs, its fresh "
                    "gradient tensors added into one fp32 bucket with torch ops; no front-end",
            "bucket_bytes": bucket_bytes, "streams": 1,
            "allreduce": "one blocking NCCL all-reduce of the bucket per step" if world > 1 else "none (1 rank)",
        }

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up, then the timed region -------------------------------------------------------------------------
    for i in range(max(args.warmup, 3)):  # the warm-up steps ARE the timed step: NCCL, allocator and caches settle here
        step(i)
    barrier()
    launches0 = C.launch_count() if args.impl == "ours" else 0
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    evs[0].record()
    for i in range(args.steps):
        step(args.warmup + i)
        evs[i + 1].record()
    torch.cuda.synchronize()
    elapsed_ms = evs[0].elapsed_time(evs[-1])
    step_ms = [evs[i].elapsed_time(evs[i + 1]) for i in range(args.steps)]
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = (C.launch_count() - launches0) if args.impl == "ours" else None
    my_ms = elapsed_ms
    elapsed_ms = sharding.all_reduce_max(elapsed_ms, dev)
    views_total = world * vps * args.steps
    value = views_total / (elapsed_ms / 1e3)
    R = int(last_R[0])
    s_bytes = 12 * K
    ss = sorted(step_ms)

    unit = "views/s"
    out = {
        "metric": METRIC if (args.config == "C2" and not fwd_only) else
                  f"raster {'forward' if fwd_only else 'fwd+bwd'} views/s @{P} Gaussians {W}x{H} ({args.config})",
        "value": round(value, 2), "unit": unit, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(elapsed_ms / args.steps, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {
            "workload": f"{args.config}-shaped synthetic scene: {P} Gaussians from {args.V} pointmap views, {W}x{H}, SH degree "
                        f"{D}, shs+scales/rotations inputs; step = {vps} views/rank, "
                        + ("forward only" if fwd_only else "fwd+bwd with a dense dL/dcolour, Gaussian-parameter gradients "
                           "summed over the rank's views into one fp32 bucket, summed across ranks once per step (N > 1)"),
            "views_per_step_per_rank": vps, "P": P, "W": W, "H": H, "sh_degree": D, "forward_only": fwd_only,
            "l2_policy": "inputs larger than L2 (per step each rank streams > %d MB of distinct per-Gaussian inputs; "
                         "L2 is 126 MB)" % ((P * (28 * vps + 44 + s_bytes)) // 1_000_000),
            "parallelism": f"views sharded over {world} rank(s), Gaussians replicated",
        },
        "impl_detail": impl_detail,
        "num_rendered": R,  # (Gaussian, tile) instances of the step's last view
        "step_ms": {"min": round(ss[0], 4), "median": round(ss[len(ss) // 2], 4), "max": round(ss[-1], 4),
                    "first5": [round(x, 3) for x in step_ms[:5]],
                    "all": [round(x, 3) for x in step_ms] if args.steps <= 64 else None},
    }
    if world > 1:
        stat = torch.tensor([my_ms / args.steps, float(sum(costs[v] for v in my_views))], dtype=torch.float64, device=dev)
        stats = [torch.zeros_like(stat) for _ in range(world)]
        dist.all_gather(stats, stat)
        out["per_rank"] = {"ms_per_step": [round(float(t[0]), 4) for t in stats],
                           "sum_num_rendered": [int(t[1]) for t in stats],
                           "views": "dealt by cost (num_rendered)" if not args.unbalanced else "by position"}
    if world > 1 and not fwd_only and args.impl == "ours" and args.sweep_slices:
        for ns in [int(x) for x in args.sweep_slices.split(",") if x]:
            holder["no_allreduce"] = ns == 0  # 0: the step without its collective (per-rank compute time, stragglers)
            holder["bucket"] = make_bucket(max(ns, 1))
            for i in range(5):
                step(i)
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(args.steps):
                step(i)
            e1.record()
            torch.cuda.synchronize()
            if reducer is not None and not holder["no_allreduce"]:  # timeline of one more step
                t0 = torch.cuda.Event(enable_timing=True)
                t1 = torch.cuda.Event(enable_timing=True)
                reducer.trace = []
                t0.record()
                step(0)
                t1.record()
                torch.cuda.synchronize()
                tl = [(round(t0.elapsed_time(a), 3), round(t0.elapsed_time(b), 3)) for a, b in reducer.trace]
                reducer.trace = None
                if rank == 0:
                    sys.stderr.write(f"[sweep] slices={ns} timeline ms from step start (slice ready, reduced): {tl} "
                                     f"step end {t0.elapsed_time(t1):.3f}\n")
            mine = torch.tensor([e0.elapsed_time(e1) / args.steps], dtype=torch.float64, device=dev)
            every = [torch.zeros_like(mine) for _ in range(world)]
            dist.all_gather(every, mine)
            ms = max(float(t) for t in every)
            if rank == 0:
                sys.stderr.write(f"[sweep] slices={ns} ms_per_step={ms:.4f} views/s={world * vps / (ms * 1e-3):.1f} "
                                 f"per_rank_ms={[round(float(t), 3) for t in every]}\n")
        holder["no_allreduce"] = False
        holder["bucket"] = bucket
    if world > 1 and not fwd_only:
        # the collective alone, on an idle GPU: what the step has to hide
        flat = bucket.flat
        for _ in range(3):
            dist.all_reduce(flat)
        barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for _ in range(10):
            dist.all_reduce(flat)
        a1.record()
        torch.cuda.synchronize()
        ar_ms = sharding.all_reduce_max(a0.elapsed_time(a1) / 10, dev)
        out["allreduce"] = {"bytes": bucket_bytes, "nccl_alone_ms": round(ar_ms, 4),
                            "nccl_algbw_gbs": round(bucket_bytes / (ar_ms * 1e-3) / 1e9, 1),
                            "nccl_kernel": "ncclDevKernel_AllReduce_Sum_f32_*"}
        if args.impl == "ours" and reducer is not None:
            # this library's in-switch kernel alone, and its result against NCCL's on the same data
            g_ = torch.Generator(device=dev).manual_seed(100 + rank)
            flat.copy_(torch.randn(flat.numel(), device=dev, generator=g_))
            want = flat.clone()
            dist.all_reduce(want)
            reducer.all_reduce(flat, 0, flat.numel())
            reducer.wait()
            torch.cuda.synchronize()
            err = float((flat - want).abs().max() / want.abs().max())
            for _ in range(3):
                reducer.all_reduce(flat, 0, flat.numel())
            reducer.wait()
            barrier()
            a0.record()
            for _ in range(10):
                reducer.all_reduce(flat, 0, flat.numel())
            reducer.wait()
            a1.record()
            torch.cuda.synchronize()
            nv_ms = sharding.all_reduce_max(a0.elapsed_time(a1) / 10, dev)
            out["allreduce"].update({"nvls_alone_ms": round(nv_ms, 4),
                                     "nvls_algbw_gbs": round(bucket_bytes / (nv_ms * 1e-3) / 1e9, 1),
                                     "nvls_kernel": "das3r::nvls_allreduce_kernel (between two symmetric-memory barriers)",
                                     "nvls_vs_nccl_max_rel_diff": err})
    if args.impl != "ours":
        out["impl"] = "reference"
        out["cpu_baseline"] = {"value": out["value"], "unit": unit, "cores": 0, "kind": "reference",
                               "sample": "reference CUDA rasterizer (oracle/_ref, sm_100 build of the vendored "
                                         "diff-gaussian-rasterization with DAS3R's 0.001 near plane) on the GPU: the "
                                         "reference has no CPU implementation of this path; the CPU port is timed by "
                                         "`--impl oracle-cpu` and in the `ours` line's cpu_baseline"}
        out["e2e"] = {"value": out["value"], "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
        out["clocks"] = clocks
        if rank == 0:
            emit((out))
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- per-stage device times (CUDA events on the launching stream), profiled pass right after the timed region ---
    C.prof_collect(True)
    C.prof_enable(True)
    prof_steps = max(2, min(args.steps, 10))
    for i in range(prof_steps):
        step(i)
    torch.cuda.synchronize()
    collected = [(name, ms, calls) for name, ms, calls in C.prof_collect(True) if calls > 0]
    stages = {name: (ms / calls) for name, ms, calls in collected}
    stage_total = {name: ms for name, ms, calls in collected}
    C.prof_enable(False)
    peak, peak_src = hbm_peak()
    Rs = [int(r) for r in all_R[0]]
    sb = stage_bytes(P, Rs, N, T, s_bytes, vps, vps, fwd_only)
    n_sl = len(bucket.slices)
    for k in ("preprocess_bwd", "frontend_bwd"):  # one launch per bucket slice
        if k in sb:
            sb[k] = sb[k] // n_sl
    kern = {k: v for k, v in stages.items() if k in sb}
    dom = max(kern, key=lambda k: stage_total[k])  # the kernel with the largest share of the step
    achieved = sb[dom] / (kern[dom] * 1e-3) / 1e9
    vb = view_bytes(P, sum(Rs) / len(Rs), N, T, s_bytes, fwd_only)
    per_view_ms = elapsed_ms / (vps * args.steps)
    step_must_move = sum(sb[k] * (n_sl if k in ("preprocess_bwd", "frontend_bwd") else 1) for k in sb)
    traffic, traffic_src = None, None
    for cand in ("r02_traffic.json", "r01_traffic.json"):
        try:  # DRAM bytes per launch from a committed `ncu --set full` capture (not measured by this run)
            t = json.load(open(os.path.join(ROOT, "profiles", cand)))
            if dom in t:
                traffic, traffic_src = t[dom], f"profiles/{cand} (ncu --set full capture; not re-measured by this run)"
                break
        except Exception:
            pass
    out["roofline"] = {
        "bound": "hbm", "kernel": dom, "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
        "frac": round(achieved / peak, 4), "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
        "algorithmic_bytes_per_launch": sb[dom], "avg_launch_ms": round(kern[dom], 5),
        "views_per_launch": vps,
        "timing": f"CUDA events around every stage on the launching stream, {prof_steps} steps right after the timed "
                  "region",
        "stage_ms": {k: round(v, 5) for k, v in stages.items()},
        "stage_share_of_step": {k: round(v / sum(stage_total.values()), 4) for k, v in stage_total.items()},
        "sort_gkeys_per_s": round(sum(Rs) / (stages["tile_sort"] * 1e-3) / 1e9, 2) if "tile_sort" in stages else None,
        "stage_frac_of_hbm_peak": {k: round(sb[k] / (v * 1e-3) / 1e9 / peak, 4) for k, v in kern.items()},
        "view": {"algorithmic_bytes": int(vb), "achieved_gbs": round(vb / (per_view_ms * 1e-3) / 1e9, 1),
                 "frac": round(vb / (per_view_ms * 1e-3) / 1e9 / peak, 4),
                 "roofline_views_per_s_per_gpu": round(peak * 1e9 / vb, 1),
                 "what": "SURVEY.md §8d per-view bytes (rasterizer only, each view on its own) x views/s"},
        "step": {"must_move_bytes": int(step_must_move),
                 "achieved_gbs": round(step_must_move / (elapsed_ms / args.steps * 1e-3) / 1e9, 1),
                 "frac": round(step_must_move / (elapsed_ms / args.steps * 1e-3) / 1e9 / peak, 4),
                 "what": "bytes the fused step (front-end included, shared arrays read once for all views) has to move"},
    }
    out["gpu_launches"] = int(launches)
    out["clocks"] = clocks

    # ---- e2e: public API (render_inputs + MultiViewRasterizer + autograd), host inputs from pinned memory -----------
    if not args.no_e2e:
        out["e2e"] = run_e2e(args, sc, dev, world, rank, S, dL, my_views, fwd_only)
    # ---- CPU baseline: the C port of the reference algorithm on the host cores (rank 0, N = 1 only) -----------------
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(args, sc)
    if rank == 0:
        emit((out))
    if world > 1:
        dist.destroy_process_group()


def run_e2e(args, sc, dev, world, rank, S, dL, my_views, fwd_only):
    """views/s through the public Python API with HOST inputs. Per step, inside the timed region: H2D (pinned) of the
    step's poses, per view of the camera bundle (viewmatrix 16 + projmatrix 16 + campos 3 + bg 3 floats) and of
    dL/d(colour) [3,H,W]; front-end; forward; loss; autograd backward down to the raw parameters and the poses; D2H of
    every rendered image and of the step's loss. The raw Gaussian parameters are the resident model state (like
    weights). For N > 1 the leaf gradients are all-reduced once per step (blocking).

    Surfaces timed: `render_inputs` + `MultiViewRasterizer` (the headline), the same with pre-posed inputs
    (`pre_posed`), and the drop-in per-view `GaussianRasterizer` (`per_view_api`)."""
    from das3r_b200.rasterizer import GaussianRasterizationSettings, GaussianRasterizer
    from das3r_b200.multiview import MultiViewRasterizer
    from das3r_b200.frontend import render_inputs
    from das3r_b200.synth import view_inputs
    from das3r_b200 import sharding

    W, H, D, vps, P = args.W, args.H, args.sh_degree, args.views_per_step, args.P
    cam_row = torch.cat([S["view"].flatten().cpu(), S["proj"].flatten().cpu(), S["campos"].cpu(), S["bg"].cpu()])
    cam_host = cam_row[None].repeat(vps, 1).contiguous().pin_memory()        # one camera bundle per view, [vps,38]
    dL_host = dL.cpu()[None].repeat(vps, 1, 1, 1).contiguous().pin_memory()  # one dL/dcolour per view, [vps,3,H,W]
    cam_dev = torch.empty_like(cam_host, device=dev)   # the step's host inputs travel as one copy per kind
    dL_dev = torch.empty_like(dL_host, device=dev)
    img_host = [torch.empty(3, H, W).pin_memory() for _ in range(vps)]
    loss_host = torch.empty(1).pin_memory()
    mv = MultiViewRasterizer()
    copy_stream = torch.cuda.Stream(device=dev)

    def settings_for(v):
        c = cam_dev[v]
        return GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=S["tanx"], tanfovy=S["tany"], bg=c[35:38], scale_modifier=1.0,
            viewmatrix=c[0:16].view(4, 4), projmatrix=c[16:32].view(4, 4), sh_degree=D, campos=c[32:35],
            prefiltered=False, debug=False)

    raw = dict(xyz=sc.xyz.to(dev), rotation=sc.rot.to(dev), opacity=torch.logit(sc.opacity.clamp(1e-6, 1 - 1e-6)).to(dev),
               scaling=torch.log(sc.scales).to(dev), features_dc=sc.shs[:, :1].contiguous().to(dev),
               features_rest=sc.shs[:, 1:].contiguous().to(dev), conf_static=torch.ones(P, device=dev))
    raw = {k: v.contiguous().requires_grad_(True) for k, v in raw.items()}
    poses_host = sc.poses[my_views].contiguous().pin_memory()
    poses_dev = torch.empty(vps, 7, device=dev, requires_grad=True)

    def finish_step(colors, leaves):
        main = torch.cuda.current_stream()
        if fwd_only:
            for v, c in enumerate(colors):
                img_host[v].copy_(c.detach(), non_blocking=True)
            main.synchronize()
            return
        main.wait_stream(copy_stream)
        loss = sum((c * dL_dev[v]).sum() for v, c in enumerate(colors))
        copy_stream.wait_stream(main)
        with torch.cuda.stream(copy_stream):  # images go home while the backward runs
            for v, c in enumerate(colors):
                img_host[v].copy_(c.detach(), non_blocking=True)
                c.record_stream(copy_stream)
        loss.backward()
        loss_host.copy_(loss.detach().view(1), non_blocking=True)
        if world > 1:  # one collective for all leaf gradients (flattened), not one per tensor
            gs = [t.grad for t in leaves if t.grad is not None]
            flat = torch.cat([g_.reshape(-1) for g_ in gs])
            dist.all_reduce(flat)
            o_ = 0
            for g_ in gs:
                g_.copy_(flat[o_:o_ + g_.numel()].view_as(g_))
                o_ += g_.numel()
        main.wait_stream(copy_stream)
        main.synchronize()  # the step's images and loss are on the host now

    def host_inputs():
        main = torch.cuda.current_stream()
        cam_dev.copy_(cam_host, non_blocking=True)
        if not fwd_only:
            copy_stream.wait_stream(main)  # the previous step is done with dL_dev
            with torch.cuda.stream(copy_stream):
                dL_dev.copy_(dL_host, non_blocking=True)  # only needed by the loss: overlaps the forwards

    def step_frontend():
        leaves = list(raw.values()) + [poses_dev]
        for t in leaves:
            t.grad = None
        with torch.no_grad():
            poses_dev.copy_(poses_host, non_blocking=True)
        host_inputs()
        ctx = torch.no_grad() if fwd_only else torch.enable_grad()
        with ctx:
            fe = render_inputs(raw["xyz"], raw["rotation"], raw["opacity"], raw["scaling"], raw["features_dc"],
                               raw["features_rest"], raw["conf_static"], None, poses_dev, sh_degree=D)
            colors, _radii = mv([settings_for(v) for v in range(vps)], fe.means3D, fe.rotations, fe.opacity, fe.shs,
                                fe.scales)
            finish_step(colors, leaves)

    # pre-posed inputs (the rasterizer alone, like the reference arm)
    pool = []
    for v in my_views:
        vi = view_inputs(sc, v)
        pool.append(dict(means=vi["means3D"].contiguous().to(dev).requires_grad_(True),
                         rots=vi["rotations"].contiguous().to(dev).requires_grad_(True)))
    shared = dict(shs=S["shs"].clone().requires_grad_(True), opac=S["opac"].clone().requires_grad_(True),
                  scales=S["scales"].clone().requires_grad_(True))
    means2D = torch.zeros(P, 3, device=dev, requires_grad=True)
    pp_leaves = list(shared.values()) + [means2D] + [x for pv in pool for x in pv.values()]

    def step_pre_posed():
        for t in pp_leaves:
            t.grad = None
        host_inputs()
        ctx = torch.no_grad() if fwd_only else torch.enable_grad()
        with ctx:
            colors, _radii = mv([settings_for(v) for v in range(vps)], [pv["means"] for pv in pool],
                                [pv["rots"] for pv in pool], shared["opac"], shared["shs"], shared["scales"])
            finish_step(colors, list(shared.values()) + [pool[0]["means"], pool[0]["rots"]])

    def step_per_view():
        for t in pp_leaves:
            t.grad = None
        total = None
        ctx = torch.no_grad() if fwd_only else torch.enable_grad()
        with ctx:
            for v, pv in enumerate(pool):
                cam_dev[v].copy_(cam_host[v], non_blocking=True)
                if not fwd_only:
                    dL_dev[v].copy_(dL_host[v], non_blocking=True)
                color, _radii = GaussianRasterizer(settings_for(v))(
                    means3D=pv["means"], means2D=means2D, opacities=shared["opac"], shs=shared["shs"],
                    scales=shared["scales"], rotations=pv["rots"])
                if not fwd_only:
                    loss = (color * dL_dev[v]).sum()
                    loss.backward()
                    total = loss.detach() if total is None else total + loss.detach()
                img_host[v].copy_(color.detach(), non_blocking=True)
        if total is not None:
            loss_host.copy_(total.view(1), non_blocking=True)
            if world > 1:
                for t in list(shared.values()) + [pool[0]["means"], pool[0]["rots"]]:
                    dist.all_reduce(t.grad)
        torch.cuda.current_stream().synchronize()

    def timed(step, steps):
        for _ in range(3):
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

    steps = max(3, min(args.steps // 3, 40))
    v_front = timed(step_frontend, steps)
    v_multi = timed(step_pre_posed, steps)
    v_single = timed(step_per_view, steps)
    img_b = 3 * H * W * 4
    return {"value": v_front, "unit": "views/s",
            "h2d_bytes_per_step": vps * (38 * 4 + (0 if fwd_only else img_b)) + vps * 7 * 4,
            "d2h_bytes_per_step": vps * img_b + (0 if fwd_only else 4),
            "steps": steps,
            "what": "das3r_b200.frontend.render_inputs + das3r_b200.multiview.MultiViewRasterizer + autograd: raw model "
                    "parameters and the step's poses -> images -> gradients of the raw parameters and the poses; per step "
                    "H2D (pinned) of the poses, per view of the camera bundle and dL/dcolour, D2H of every rendered image "
                    "and of the loss; Gaussian parameters resident (model state)",
            "pre_posed": {"value": v_multi, "unit": "views/s",
                          "what": "MultiViewRasterizer alone on pre-posed means/rotations (what the reference arm times), "
                                  "same copies"},
            "per_view_api": {"value": v_single, "unit": "views/s",
                             "what": "the drop-in GaussianRasterizer.forward + autograd backward called once per view, "
                                     "same copies"}}


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
        if not args.forward_only:
            oracle.backward(fw, dL, **vi)
        t += time.perf_counter() - a
    return {"value": round(n / t, 4), "unit": "views/s", "cores": os.cpu_count(), "kind": "port",
            "sample": f"{n} views ({'forward' if args.forward_only else 'fwd+bwd'}) of the same {args.P}-Gaussian "
                      f"{args.W}x{args.H} scene through the C port of the reference algorithm (oracle/raster_oracle.c, "
                      f"OpenMP on all {os.cpu_count()} host cores)"}


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
