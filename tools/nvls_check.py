"""Developer tool (2+ GPUs): das3r_nvls_allreduce_sum against NCCL and torch's symmetric-memory all-reduce."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
from das3r_b200 import load_native

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
C = load_native()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 60_000_064
flat = symm.empty(n, dtype=torch.float32, device=dev)
h = symm.rendezvous(flat, dist.group.WORLD)
if rank == 0:
    print("multicast", h.has_multicast_support, hex(h.multicast_ptr), "offset", h.offset, "buffer_size", h.buffer_size,
          "storage_offset", flat.storage_offset(), "data_ptr", hex(flat.data_ptr()), "buf0", hex(h.buffer_ptrs[rank]))
g = torch.Generator(device=dev).manual_seed(100 + rank)
src = torch.randn(n, device=dev, generator=g)
want = src.clone(); dist.all_reduce(want)
mc = int(h.multicast_ptr) + int(h.offset)
def ours(ctas):
    h.barrier(channel=0)
    C.nvls_allreduce_sum(mc, 0, n, rank, world, ctas, lr)
    h.barrier(channel=0)
flat.copy_(src); torch.cuda.synchronize(); dist.barrier()
ours(24); torch.cuda.synchronize()
d = (flat - want).abs()
half = n // world
print(rank, "ours max abs diff", float(d.max()), "per share", [float(d[i*half:(i+1)*half].max()) for i in range(world)],
      "equal-to-src frac", float((flat == src).float().mean()))
flat.copy_(src); torch.cuda.synchronize(); dist.barrier()
torch.ops.symm_mem.multimem_all_reduce_(flat, "sum", dist.group.WORLD.group_name); torch.cuda.synchronize()
print(rank, "torch multimem diff", float((flat - want).abs().max()))
def timeit(fn, it=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(it): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / it
for ctas in (16, 32, 48, 64, 96, 148):
    t = timeit(lambda: ours(ctas))
    if rank == 0: print("ours ctas", ctas, round(t, 4), "ms", round(n * 4 / t / 1e6, 1), "GB/s")
t = timeit(lambda: (h.barrier(channel=0), h.barrier(channel=0)))
if rank == 0: print("two barriers alone", round(t, 4), "ms")
for frac in (2, 4, 8):
    m = (n // frac) // 16 * 16
    def part():
        h.barrier(channel=0); C.nvls_allreduce_sum(mc, 0, m, rank, world, 16, lr); h.barrier(channel=0)
    t = timeit(part)
    if rank == 0: print(f"ours 1/{frac} of the buffer, 16 ctas", round(t, 4), "ms")
    def part32():
        h.barrier(channel=0); C.nvls_allreduce_sum(mc, 0, m, rank, world, 48, lr); h.barrier(channel=0)
    t = timeit(part32)
    if rank == 0: print(f"ours 1/{frac} of the buffer, 48 ctas", round(t, 4), "ms")
t = timeit(lambda: torch.ops.symm_mem.multimem_all_reduce_(flat, "sum", dist.group.WORLD.group_name))
if rank == 0: print("torch multimem_all_reduce_", round(t, 4), "ms", round(n * 4 / t / 1e6, 1), "GB/s")
t = timeit(lambda: torch.ops.symm_mem.two_shot_all_reduce_(flat, "sum", dist.group.WORLD.group_name))
if rank == 0: print("torch two_shot_all_reduce_", round(t, 4), "ms")
t = timeit(lambda: dist.all_reduce(want))
if rank == 0: print("nccl", round(t, 4), "ms", round(n * 4 / t / 1e6, 1), "GB/s")
dist.destroy_process_group()
