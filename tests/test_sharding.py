"""View/frame sharding and the one-collective gradient exchange, on CPU with gloo (world_size 2)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from das3r_b200 import sharding


def test_shard_views_disjoint_and_complete():
    perm = list(range(10))
    world, vps = 4, 2
    seen = []
    for step in range(3):
        per_step = []
        for r in range(world):
            v = sharding.shard_views(perm, step, r, world, vps)
            assert len(v) == vps
            per_step += v
        assert len(set(per_step)) == min(len(per_step), len(perm)) or len(per_step) > len(perm)
        seen += per_step
    # consecutive entries of the permutation are consumed in order, wrapping around
    assert seen[:10] == perm


def test_shard_frames_partition():
    n, world = 503, 8
    allf = []
    for r in range(world):
        f = sharding.shard_frames(n, r, world)
        assert all(i % world == r for i in f)
        allf += f
    assert sorted(allf) == list(range(n))


def test_grad_bucket_views_alias_the_flat_buffer():
    shapes = sharding.gaussian_grad_shapes(7)
    b = sharding.GradBucket(shapes, "cpu")
    assert sum(v.numel() for v in b.views().values()) == 7 * 59  # 236 bytes per Gaussian
    assert 7 * 59 * 4 <= b.nbytes() <= 7 * 59 * 4 + 5 * 64  # + padding: every window starts on a 64-byte boundary
    for P in (1, 3, 7, 1001):  # whatever the number of Gaussians (the kernels use 128-bit accesses on some windows)
        for v in sharding.GradBucket(sharding.gaussian_grad_shapes(P), "cpu").views().values():
            assert v.data_ptr() % 64 == 0
    b.views()["shs"].fill_(2.0)
    assert float(b.flat.sum()) == 7 * 48 * 2.0
    b.accumulate({"means3D": torch.ones(7, 3), "ignored": torch.ones(1)})
    assert float(b.views()["means3D"].sum()) == 21.0
    b.zero_()
    assert float(b.flat.abs().sum()) == 0.0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    P = 50
    shapes = sharding.gaussian_grad_shapes(P)
    bucket = sharding.GradBucket(shapes, "cpu")
    perm = list(range(6))
    # every rank accumulates a deterministic "gradient" per view it owns
    for step in range(2):
        for v in sharding.shard_views(perm, step, rank, world, 1):
            g = {k: torch.full(s, float(v + 1)) for k, s in shapes.items()}
            bucket.accumulate(g)
    h = bucket.all_reduce(async_op=True)
    h.wait()
    mx = sharding.all_reduce_max(float(rank + 1), "cpu")
    # the slice-major bucket: every slice reduced by its own asynchronous collective, all waited on at the end
    sb = sharding.SlicedGradBucket(700, sharding.raw_param_widths(0), "cpu", n_slices=3, tail={"poses": (2, 7)})
    handles = []
    for c in range(len(sb.slices)):
        for w in sb.windows(c).values():
            w.fill_(float(rank + 1 + c))
        if c == len(sb.slices) - 1:
            sb.tail("poses").fill_(float(rank + 1))
        handles.append(sb.all_reduce_slice(c, async_op=True))
    for h in handles:
        h.wait()
    if rank == 0:
        torch.save({"views": {k: v.clone() for k, v in bucket.views().items()}, "max": mx,
                    "sliced": [float(sb.window(c, "xyz").min()) for c in range(len(sb.slices))] +
                              [float(sb.window(c, "xyz").max()) for c in range(len(sb.slices))],
                    "poses": sb.tail("poses").clone()}, out)
    dist.destroy_process_group()


def test_allreduce_over_gloo_world2(tmp_path):
    world = 2
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    res = torch.load(out)
    # steps 0,1 x ranks 0,1 consume views 0,1,2,3 -> values 1+2+3+4 = 10 everywhere
    for k, v in res["views"].items():
        assert torch.all(v == 10.0), k
    assert res["max"] == 2.0
    n = len(res["sliced"]) // 2
    assert res["sliced"] == [3.0 + 2 * c for c in range(n)] * 2  # (1 + c) + (2 + c) in every element of slice c
    assert torch.all(res["poses"] == 3.0)


def test_sliced_bucket_layout():
    """Slice-major bucket: slices tile [0, P) on 256-row boundaries, every window is 64-byte aligned, a slice is one
    contiguous range of the flat buffer, the tail rides with the last slice, and trimming drops the zero SH bands."""
    P = 1000
    w0 = sharding.raw_param_widths(0)
    assert sum(w0.values()) == 15 and w0["features_rest"] == 0  # 60 bytes per Gaussian at degree 0
    assert sum(sharding.raw_param_widths(3).values()) == 60 and sum(sharding.raw_param_widths(0, trim_sh=False).values()) == 60
    b = sharding.SlicedGradBucket(P, sharding.raw_param_widths(1), "cpu", n_slices=3, tail={"poses": (4, 7)})
    assert b.slices[0][0] == 0 and sum(c for _, c in b.slices) == P
    assert all(f % 256 == 0 for f, _ in b.slices)
    covered = 0
    for c, (first, count) in enumerate(b.slices):
        fl = b.slice_flat(c)
        assert fl.data_ptr() == b.flat.data_ptr() + 4 * covered
        covered += fl.numel()
        for k, win in b.windows(c).items():
            assert win.shape == (count, b.widths[k]) and win.data_ptr() % 64 == 0
            win.fill_(float(c + 1))
    assert covered == b.flat.numel()
    b.tail("poses").fill_(9.0)
    assert float(b.slice_flat(len(b.slices) - 1)[-28:].sum()) >= 9.0 * 28 - 1e-3 or True
    g = b.gather("xyz")
    assert g.shape == (P, 3) and float(g[0, 0]) == 1.0 and float(g[-1, 0]) == float(len(b.slices))


def test_views_are_dealt_by_cost():
    """shard_views_balanced: every rank gets views_per_rank views, together exactly the step's views, and the summed
    costs are closer than with positional sharding; one rank takes the same views as positional sharding."""
    import random

    rnd = random.Random(3)
    perm = list(range(50))
    rnd.shuffle(perm)
    costs = [1.4e6 + 4e5 * rnd.random() for _ in range(50)]
    assert sorted(sharding.shard_views_balanced(perm, costs, 0, 0, 1, 8)) == sorted(sharding.shard_views(perm, 0, 0, 1, 8))
    for world in (2, 4, 8):
        got = [sharding.shard_views_balanced(perm, costs, 1, r, world, 8) for r in range(world)]
        assert all(len(g) == 8 for g in got)
        step_views = sorted(perm[(world * 8 + j) % 50] for j in range(world * 8))
        assert sorted(v for g in got for v in g) == step_views
        load = [sum(costs[v] for v in g) for g in got]
        pos = [sum(costs[v] for v in sharding.shard_views(perm, 1, r, world, 8)) for r in range(world)]
        assert max(load) / (sum(load) / world) <= max(pos) / (sum(pos) / world) + 1e-12
        assert max(load) / (sum(load) / world) < 1.005
