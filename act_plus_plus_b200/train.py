"""One optimisation step on the engine: zero_grad -> forward(+loss) -> backward -> AdamW, optionally data-parallel
(batch sharded across ranks, gradients averaged with NCCL over NVLink while later backward stages still run)."""
from __future__ import annotations

import os

import torch

from .dp import GradReducer
from .engine import R_BACKBONE, R_CVAE, R_DEAD, R_MAIN, ActEngine


def backward_data_parallel(eng: ActEngine, group, loss_grad: float = 1.0, bucket_elems: int = 0):
    """Staged backward: as soon as a parameter region's gradients are final its all-reduce is enqueued (NCCL runs on
    its own stream, ordered after the compute stream at enqueue time), so the transfer of the transformer gradients
    overlaps the CVAE-encoder and backbone backward.  Regions are contiguous slices of the flat gradient buffer."""
    red = GradReducer(eng.flat_grads, group, bucket_elems)
    # stage 4 = CVAE-encoder and backbone backward as two concurrent branches (ACT_DP_STAGES=123 keeps them serial,
    # which starts the CVAE region's all-reduce earlier at the price of the lost overlap)
    plan = (((1, (R_MAIN,)), (2, (R_CVAE,)), (3, (R_BACKBONE,))) if os.environ.get("ACT_DP_STAGES", "14") == "123"
            else ((1, (R_MAIN,)), (4, (R_CVAE, R_BACKBONE))))
    for stage, regions in plan:
        eng.backward(stage, loss_grad)
        for region in regions:
            off, n = eng.region(region)
            red.reduce(off, n)
    red.wait()


def backward_and_step_data_parallel(eng: ActEngine, group, loss_grad: float = 1.0, bucket_elems: int = 0):
    """Data-parallel loss.backward() + optimizer.step(): backward stage 1 (heads + transformer) -> all-reduce of that
    region -> its AdamW on a side stream as soon as the reduction lands, all of it under backward stage 4 (CVAE encoder
    || backbones); the two remaining regions are reduced and updated after stage 4."""
    cur = torch.cuda.current_stream()
    side = _side_stream(eng)
    step = eng.begin_optimizer_step()
    red_main = GradReducer(eng.flat_grads, group, bucket_elems)
    eng.backward(1, loss_grad)
    red_main.reduce(*eng.region(R_MAIN))
    side.wait_stream(cur)                       # every read of the main-region weights (stage 1) is ordered before
    with torch.cuda.stream(side):
        red_main.wait()
        eng.adamw_region(step, R_MAIN, 3)
        eng.adamw_region(step, R_DEAD, 3)
    eng.backward(4, loss_grad)
    red = GradReducer(eng.flat_grads, group, bucket_elems)
    for region in (R_CVAE, R_BACKBONE):
        red.reduce(*eng.region(region))
    red.wait()
    eng.adamw_region(step, R_CVAE)
    eng.adamw_region(step, R_BACKBONE)
    cur.wait_stream(side)
    eng.end_optimizer_step()


def backward_and_step_hooked(eng: ActEngine, group, loss_grad: float = 1.0):
    """Data-parallel loss.backward() + optimizer.step() through the engine's gradient-ready hook
    (act_engine_backward_and_step_dp): the engine announces each contiguous gradient range on the internal stream where
    it becomes final -- main region after the transformer backward, each CVAE-encoder layer as its backward ends, the
    backbone region after its weight-gradient fold -- and this callback enqueues the NCCL average there.  Every range's
    AdamW follows its own all-reduce on that stream, so collectives and updates run under the remaining backward
    branches; only the last range's reduction is exposed."""
    import torch.distributed as dist
    flat = eng.flat_grads
    world = dist.get_world_size(group)
    native_avg = dist.get_backend(group) == "nccl"

    def ready(off: int, n: int, stream_ptr: int):
        s = _external_stream(stream_ptr, flat.device)
        with torch.cuda.stream(s):
            t = flat[off:off + n]
            work = dist.all_reduce(t, op=dist.ReduceOp.AVG if native_avg else dist.ReduceOp.SUM, group=group, async_op=True)
            work.wait()                          # orders `s` after the collective; does not block the host
            if not native_avg:
                t.mul_(1.0 / world)

    eng.backward_and_step_dp(ready, loss_grad)


_EXT_STREAMS = {}


def _external_stream(ptr: int, device) -> "torch.cuda.Stream":
    key = (ptr, device.index)
    if key not in _EXT_STREAMS:
        _EXT_STREAMS[key] = torch.cuda.ExternalStream(ptr, device=device) if ptr else torch.cuda.default_stream(device)
    return _EXT_STREAMS[key]


_SIDE_STREAMS = {}


def _side_stream(eng: ActEngine) -> "torch.cuda.Stream":
    key = (id(eng), torch.cuda.current_device())
    if key not in _SIDE_STREAMS:
        _SIDE_STREAMS[key] = torch.cuda.Stream()
    return _SIDE_STREAMS[key]


def train_step(eng: ActEngine, qpos, image, actions, is_pad, eps=None, dp_group=None) -> torch.Tensor:
    """Returns the device tensor [l1, kl, loss] of this rank's shard."""
    eng.zero_grad(overlap_forward=True)
    out = eng.forward(qpos, image, actions, is_pad, eps)
    if dp_group is None:
        eng.backward_and_step(1.0)
    elif eng.dp_connected:
        # gradient exchange fused into the optimiser kernel over NVLink peer memory (eng.dp_connect(group) was called)
        eng.backward_and_step_p2p(1.0)
    elif os.environ.get("ACT_DP_HOOK", "1") == "1":
        backward_and_step_hooked(eng, dp_group, 1.0)
    elif os.environ.get("ACT_DP_OVERLAP_OPT", "1") == "1":
        backward_and_step_data_parallel(eng, dp_group, 1.0)
    else:
        backward_data_parallel(eng, dp_group, 1.0)
        eng.adamw_step()
    return out["losses"]


class InputPrefetcher:
    """Double-buffered host->device staging for the training loop (the role of the DataLoader's pinned-memory prefetch in
    imitate_episodes.py:596-601): `submit(host_batch)` enqueues the H2D copy of a pinned batch on a copy stream,
    `next()` hands the oldest submitted batch to the current stream.  Submitting batch i+1 before consuming batch i lets
    its transfer run under step i's kernels.  A slot is reused only after the step that read it has been enqueued
    (event recorded by `next()` for the previous occupant and waited for by the copy stream)."""

    def __init__(self, example: dict, device, depth: int = 2):
        self.device = torch.device(device)
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.slots = [{k: torch.empty(v.shape, dtype=v.dtype, device=self.device) for k, v in example.items()}
                      for _ in range(depth)]
        self.ready = [torch.cuda.Event() for _ in range(depth)]
        self.free = [None] * depth
        self.head = self.tail = 0
        self.depth = depth

    def submit(self, host_batch: dict):
        assert self.tail - self.head < self.depth, "prefetch queue full: call next() first"
        s = self.tail % self.depth
        with torch.cuda.stream(self.copy_stream):
            if self.free[s] is not None:
                self.copy_stream.wait_event(self.free[s])
            for k, v in host_batch.items():
                self.slots[s][k].copy_(v, non_blocking=True)
            self.ready[s].record(self.copy_stream)
        self.tail += 1

    def next(self) -> dict:
        """The returned tensors are valid until `depth` more batches have been submitted; consume them on the current
        stream (e.g. copy into the step's static inputs) before calling submit() again for this slot."""
        assert self.head < self.tail, "nothing submitted"
        s = self.head % self.depth
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(self.ready[s])
        self.head += 1
        return self.slots[s]

    def consumed(self):
        """Mark the most recently returned batch as read by everything enqueued so far on the current stream."""
        s = (self.head - 1) % self.depth
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.free[s] = ev
