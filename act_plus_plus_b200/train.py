"""One optimisation step on the engine: zero_grad -> forward(+loss) -> backward -> AdamW, optionally data-parallel
(batch sharded across ranks, gradients averaged with NCCL over NVLink while later backward stages still run)."""
from __future__ import annotations

import torch

from .dp import GradReducer
from .engine import R_BACKBONE, R_CVAE, R_MAIN, ActEngine


def backward_data_parallel(eng: ActEngine, group, loss_grad: float = 1.0, bucket_elems: int = 0):
    """Staged backward: as soon as a parameter region's gradients are final its all-reduce is enqueued (NCCL runs on
    its own stream, ordered after the compute stream at enqueue time), so the transfer of the transformer gradients
    overlaps the CVAE-encoder and backbone backward.  Regions are contiguous slices of the flat gradient buffer."""
    red = GradReducer(eng.flat_grads, group, bucket_elems)
    # (used when loss.backward() and optimizer.step() stay separate calls; the fused step goes through
    # backward_and_step_hooked / the peer-memory kernel)  stage 4 = CVAE-encoder and backbone backward as two branches
    for stage, regions in ((1, (R_MAIN,)), (4, (R_CVAE, R_BACKBONE))):
        eng.backward(stage, loss_grad)
        for region in regions:
            off, n = eng.region(region)
            red.reduce(off, n)
    red.wait()


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


def train_step(eng: ActEngine, qpos, image, actions, is_pad, eps=None, dp_group=None, advance_on_device: bool = False) -> torch.Tensor:
    """One optimisation step; returns the device tensor [l1, kl, loss] of this rank's shard.  The gradient clear runs
    beside the forward, every parameter region's AdamW beside the remaining backward.  Data parallel: the exchange is
    fused into the optimiser kernel over NVLink peer memory when `eng.dp_connect(group)` succeeded, NCCL all-reduces
    enqueued from the engine's gradient-ready hook otherwise.  advance_on_device: see ActEngine.backward_and_step."""
    eng.zero_grad(overlap_forward=True)
    out = eng.forward(qpos, image, actions, is_pad, eps)
    if dp_group is None:
        eng.backward_and_step(1.0, advance_on_device=advance_on_device)
    elif eng.dp_connected:
        eng.backward_and_step_p2p(1.0, advance_on_device=advance_on_device)
    else:
        backward_and_step_hooked(eng, dp_group, 1.0)
    return out["losses"]


class GraphedTrainStep:
    """`train_step` at fixed shapes as ONE CUDA graph (no tracing compiler: the engine's own launch sequence, with its
    forked streams turned into parallel graph branches).  `step(batch)` copies the batch into static device buffers and
    replays; the optimiser step number and the dropout step advance through device counters inside the graph, so every
    replay is a genuine new step.  Falls back to direct launches when capture is not possible."""

    def __init__(self, eng: ActEngine, example: dict, dp_group=None, warmup: int = 2, stream_priority: int = -1):
        self.eng, self.group = eng, dp_group
        self.static = {k: torch.empty_like(v, device=eng.device) for k, v in example.items()}
        self.graph, self.losses = None, None
        for k, v in example.items():
            self.static[k].copy_(v)
        for _ in range(max(1, warmup)):                      # eager steps: lazy host set-up, device counters seeded
            self._direct(self.static)
        torch.cuda.synchronize(eng.device)
        native = dp_group is None or eng.dp_connected        # the NCCL-hook path counts its step on the host
        try:
            s = torch.cuda.Stream(device=eng.device, priority=stream_priority)
            s.wait_stream(torch.cuda.current_stream(eng.device))
            with torch.cuda.stream(s):
                self._direct(self.static)
            torch.cuda.current_stream(eng.device).wait_stream(s)
            torch.cuda.synchronize(eng.device)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self.losses = train_step(eng, self.static["qpos"], self.static["image"], self.static["actions"],
                                         self.static["is_pad"], self.static["eps"], dp_group, advance_on_device=native)
            self.graph, self._native = g, native
            torch.cuda.synchronize(eng.device)
        except Exception as ex:                              # capture is an optimisation, never a requirement
            import warnings
            warnings.warn(f"act_plus_plus_b200: CUDA-graph capture of the training step failed ({ex}); launching directly")
            self.graph = None

    def _direct(self, b):
        return train_step(self.eng, b["qpos"], b["image"], b["actions"], b["is_pad"], b["eps"], self.group)

    def step(self, batch: dict) -> torch.Tensor:
        if self.graph is None:
            return self._direct(batch)
        for k, v in batch.items():
            self.static[k].copy_(v, non_blocking=True)       # device->device (or pinned host->device), part of the step
        self.graph.replay()
        if self._native:
            self.eng.note_graph_step()
        else:
            self.eng._shared["version"] += 1; self.eng._synced_version = self.eng._shared["version"]
        return self.losses


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
