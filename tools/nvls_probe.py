"""Probe of the NVSwitch multicast (NVLS) mapping the data-parallel exchange uses (run under torchrun, one rank per GPU):
symmetric memory from torch.distributed._symmetric_memory, then the package's own multimem kernels through the C-ABI.
  1. every rank fills its buffer with rank+1; act_multimem_reduce_f32 on the multicast address must return world*(world+1)/2
  2. rank r broadcasts the value 100+r into slice r with act_multimem_bcast_f32; after a barrier every rank must see all slices
  3. bandwidth of both over 64 MB"""
import os
import sys
import torch
import torch.distributed as dist
sys.path.insert(0, ".")
from act_plus_plus_b200 import _lib

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=dev)
lib = _lib.lib()
import torch.distributed._symmetric_memory as symm
N = 16 << 20                                         # 64 MB of fp32
try:
    buf = symm.empty(N, dtype=torch.float32, device=dev)
    h = symm.rendezvous(buf, dist.group.WORLD)
    mc = int(h.multicast_ptr)
    print(f"rank {rank}: multicast_ptr {mc:#x} buffer_ptrs {[hex(int(p)) for p in h.buffer_ptrs]} local {buf.data_ptr():#x}", flush=True)
except Exception as ex:
    print(f"rank {rank}: symmetric memory unavailable: {type(ex).__name__}: {ex}", flush=True)
    mc = 0
ok = torch.tensor([1.0 if mc else 0.0], device=dev)
dist.all_reduce(ok, op=dist.ReduceOp.MIN)
if float(ok) == 0.0:
    if rank == 0:
        print("nvls_probe: NO multicast support", flush=True)
    dist.destroy_process_group()
    sys.exit(0)
st = torch.cuda.current_stream().cuda_stream
out = torch.empty(N, dtype=torch.float32, device=dev)
buf.fill_(float(rank + 1))
torch.cuda.synchronize(); dist.barrier()
assert lib.act_multimem_reduce_f32(mc, out.data_ptr(), N, st) == 0, _lib.last_error()
torch.cuda.synchronize()
want = world * (world + 1) / 2
good1 = bool((out == want).all())
dist.barrier()
per = N // world
src = torch.full((per,), 100.0 + rank, dtype=torch.float32, device=dev)
assert lib.act_multimem_bcast_f32(mc + 4 * per * rank, src.data_ptr(), per, st) == 0, _lib.last_error()
torch.cuda.synchronize(); dist.barrier()
good2 = all(bool((buf[r * per:(r + 1) * per] == 100.0 + r).all()) for r in range(world))
print(f"rank {rank}: ld_reduce {'OK' if good1 else 'WRONG'} (got {float(out[0])}, want {want}); multimem.st {'OK' if good2 else 'WRONG'}", flush=True)
# bandwidth: every rank reduces / broadcasts its own 1/world slice at the same time (the exchange's access pattern)
for name, fn, args in (("ld_reduce", lib.act_multimem_reduce_f32, (mc + 4 * per * rank, out.data_ptr(), per, st)),
                       ("st", lib.act_multimem_bcast_f32, (mc + 4 * per * rank, src.data_ptr(), per, st))):
    for _ in range(3):
        fn(*args)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        fn(*args)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(f"rank {rank}: multimem.{name} of a {per * 4 / 1e6:.1f} MB slice on every rank at once: {ms * 1e3:.1f} us = {per * 4 / ms / 1e6:.1f} GB/s per rank", flush=True)
    dist.barrier()
ok = torch.tensor([1.0 if good1 and good2 else 0.0], device=dev)
dist.all_reduce(ok, op=dist.ReduceOp.MIN)
if rank == 0:
    print("nvls_probe:", "PASS" if float(ok) == 1.0 else "FAIL", flush=True)
dist.destroy_process_group()
