"""das3r_nvls_allreduce_sum (include/das3r_comm.h) through sharding.NvlsReducer on 2 GPUs of one NVSwitch box against
NCCL's all-reduce of the same data: slices of a slice-major bucket reduced one after the other, asynchronously, on the
reducer's stream. Needs 2 GPUs with multicast support; skipped elsewhere (the 1-GPU test box)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from das3r_b200 import sharding

    red = sharding.NvlsReducer(dev)
    b = sharding.SlicedGradBucket(100_003, sharding.raw_param_widths(1), dev, n_slices=3, tail={"poses": (4, 7)},
                                  allocate=red.allocate)
    res = {"available": red.available}
    if red.available:
        g = torch.Generator(device=dev).manual_seed(7 + rank)
        b.flat.copy_(torch.randn(b.flat.numel(), device=dev, generator=g))
        want = b.flat.clone()
        dist.all_reduce(want)
        for c in range(len(b.slices)):
            red.all_reduce(b.flat, *b.slice_bounds(c))
        red.wait()
        torch.cuda.synchronize()
        res["err"] = float((b.flat - want).abs().max() / want.abs().max())
        # a second round on the reduced data (what the next step's bucket would see): still consistent across ranks
        for c in range(len(b.slices)):
            red.all_reduce(b.flat, *b.slice_bounds(c))
        red.wait()
        torch.cuda.synchronize()
        res["err2"] = float((b.flat - world * want).abs().max() / want.abs().max())
    if rank == 0:
        torch.save(res, out)
    dist.barrier()
    dist.destroy_process_group()


def test_in_switch_allreduce_matches_nccl(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs on one NVSwitch box")
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    res = torch.load(out)
    if not res["available"]:
        pytest.skip("no NVSwitch multicast support on this box")
    assert res["err"] <= 1e-6 and res["err2"] <= 1e-6, res
