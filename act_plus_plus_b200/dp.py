"""Data-parallel plumbing for the ACT hot path (new work: the reference is single-GPU).

Training: the batch is sharded across ranks (equal shards), parameters are replicated, and the flat fp32
gradient buffer is averaged region by region -- each region's all-reduce is enqueued as soon as the staged
backward has finished it, so the NCCL transfer over NVLink overlaps the remaining backward stages.
Rollout inference: environments are sharded statically; there is no collective on the data path.

Everything here works on any torch.distributed backend (NCCL on the GPUs; gloo in the CPU tests)."""
from __future__ import annotations

from typing import List, Sequence, Tuple

import torch


def shard_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous [start, stop) of `n` items owned by `rank`; the remainder goes to the lowest ranks."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, rem = divmod(n, world)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def owner_slices(count: int, world: int) -> List[Tuple[int, int, int]]:
    """(rank, lo, hi): the floats [lo, hi) of a `count`-float range (relative to its start) that `rank` owns in the
    peer-memory optimiser step (csrc/dp_p2p.cu::dp_slice: float4 granularity, remainder to the lowest ranks; the
    count % 4 tail belongs to rank 0)."""
    n4 = count // 4
    out = []
    for r in range(world):
        lo4, hi4 = shard_range(n4, r, world)
        if hi4 > lo4:
            out.append((r, 4 * lo4, 4 * hi4))
    if count % 4:
        out.append((0, 4 * n4, count))
    return out


def split_buckets(offset: int, count: int, bucket_elems: int) -> List[Tuple[int, int]]:
    """Cut one contiguous region into buckets of at most `bucket_elems` elements (the last may be short)."""
    if bucket_elems <= 0 or count <= bucket_elems:
        return [(offset, count)] if count > 0 else []
    out, o = [], offset
    while o < offset + count:
        n = min(bucket_elems, offset + count - o)
        out.append((o, n))
        o += n
    return out


class GradReducer:
    """Averages slices of one flat gradient tensor across a process group, asynchronously.

    reduce(off, n) enqueues the all-reduce of flat[off:off+n] (optionally cut into buckets); wait() blocks the
    current stream / thread until every enqueued reduction has finished and the result is the mean over ranks."""

    def __init__(self, flat: torch.Tensor, group=None, bucket_elems: int = 0):
        import torch.distributed as dist
        self.flat, self.group, self.bucket_elems = flat, group, bucket_elems
        self.world = dist.get_world_size(group)
        backend = dist.get_backend(group)
        self._native_avg = backend == "nccl"          # gloo has no ReduceOp.AVG: sum, then scale in wait()
        self._pending: List[tuple] = []

    def reduce(self, offset: int, count: int):
        import torch.distributed as dist
        op = dist.ReduceOp.AVG if self._native_avg else dist.ReduceOp.SUM
        for off, n in split_buckets(offset, count, self.bucket_elems):
            t = self.flat[off:off + n]
            self._pending.append((dist.all_reduce(t, op=op, group=self.group, async_op=True), t))

    def wait(self):
        for work, t in self._pending:
            work.wait()
            if not self._native_avg:
                t.mul_(1.0 / self.world)
        self._pending.clear()


def average_scalars(values: torch.Tensor, group=None) -> torch.Tensor:
    """Mean over ranks of a small tensor of per-rank means (loss logging only; equal shard sizes)."""
    import torch.distributed as dist
    out = values.clone()
    dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)
    return out / dist.get_world_size(group)


def gather_env_actions(local: torch.Tensor, counts: Sequence[int], group=None) -> torch.Tensor:
    """Rank-ordered concatenation of per-rank [E_r, A] action blocks (host-side collection of rollout actions)."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    emax = max(counts)
    pad = torch.zeros((emax,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[:local.shape[0]] = local
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad, group=group)
    return torch.cat([b[:c] for b, c in zip(bufs, counts)], dim=0)
