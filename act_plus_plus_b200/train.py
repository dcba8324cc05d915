"""One optimisation step on the engine: zero_grad -> forward(+loss) -> backward -> AdamW, optionally data-parallel
(batch sharded across ranks, gradients averaged with NCCL over NVLink while later backward stages still run)."""
from __future__ import annotations

import os

import torch

from .dp import GradReducer
from .engine import R_BACKBONE, R_CVAE, R_MAIN, ActEngine


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


def train_step(eng: ActEngine, qpos, image, actions, is_pad, eps=None, dp_group=None) -> torch.Tensor:
    """Returns the device tensor [l1, kl, loss] of this rank's shard."""
    eng.zero_grad(overlap_forward=True)
    out = eng.forward(qpos, image, actions, is_pad, eps)
    if dp_group is None:
        eng.backward_and_step(1.0)
    else:
        backward_data_parallel(eng, dp_group, 1.0)
        eng.adamw_step()
    return out["losses"]
