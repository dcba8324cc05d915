"""Batched rollout inference with the eval-time temporal ensemble (imitate_episodes.py:307-310, 338-339, 368-411,
436-438) for E environments at once, environments sharded statically across ranks.

The reference keeps one `all_time_actions[T, T+nq, A]` buffer per episode and runs six eager ops plus a host numpy
round trip per step; here every environment owns a ring `[nq, nq, A]` on its GPU and one kernel
(`act_temporal_agg_step`) does scatter + gather + `!= 0` filter + float64 exponential weighting for all of them."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib
from ._lib import check
from .dp import shard_range


class TemporalEnsembler:
    """E independent ensembles.  step(chunks[E,nq,A] f32, t) -> raw_action [E,A] float64 (reference dtype, :410-411)."""

    def __init__(self, num_envs: int, num_queries: int, action_dim: int, k: float = 0.01, device=None):
        if not torch.cuda.is_available():
            raise _lib.ActB200Error("TemporalEnsembler needs a CUDA device; there is no CPU fallback")
        self.E, self.nq, self.A, self.k = num_envs, num_queries, action_dim, float(k)
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self.ring = torch.zeros((num_envs, num_queries, num_queries, action_dim), dtype=torch.float32, device=self.device)
        self.out = torch.empty((num_envs, action_dim), dtype=torch.float64, device=self.device)
        self.t = torch.zeros(num_envs, dtype=torch.int32, device=self.device)
        self.lib = _lib.lib()

    def reset(self, env_mask: Optional[torch.Tensor] = None):
        """New episode (imitate_episodes.py:338-339 re-allocates the buffer): clear rings and step counters."""
        if env_mask is None:
            self.ring.zero_(); self.t.zero_()
        else:
            self.ring[env_mask] = 0; self.t[env_mask] = 0

    def step(self, chunks: torch.Tensor, t=None) -> torch.Tensor:
        assert chunks.is_cuda and chunks.dtype == torch.float32 and tuple(chunks.shape) == (self.E, self.nq, self.A)
        chunks = chunks.contiguous()
        if t is not None:
            if isinstance(t, int):
                self.t.fill_(t)
            else:
                self.t.copy_(t.to(torch.int32))
        check(self.lib.act_temporal_agg_step(self.ring.data_ptr(), chunks.data_ptr(), self.t.data_ptr(),
                                             self.out.data_ptr(), self.E, self.nq, self.A, C.c_double(self.k),
                                             torch.cuda.current_stream().cuda_stream), "act_temporal_agg_step")
        if t is None:
            self.t += 1
        return self.out


class BatchedRollout:
    """One policy query per environment per step (query_frequency = 1 with temporal_agg, imitate_episodes.py:308-309)
    followed by the ensemble; optional device-side pre/post-processing (qpos z-score :368-371, action
    de-normalisation :436-438) so the only host traffic per step is observations in, [E, A] actions out."""

    def __init__(self, policy, num_envs_global: int, rank: int = 0, world: int = 1, stats: Optional[dict] = None,
                 temporal_agg: bool = True):
        self.policy = policy
        self.lo, self.hi = shard_range(num_envs_global, rank, world)
        self.E = self.hi - self.lo
        m = policy.model
        self.nq, self.A = m.num_queries, m.action_dim
        self.temporal_agg = temporal_agg
        self.ens = TemporalEnsembler(self.E, self.nq, self.A, device=m.device) if temporal_agg else None
        self.stats = None
        if stats is not None:
            dev = m.device
            self.stats = {k: torch.as_tensor(v, dtype=torch.float32, device=dev) for k, v in stats.items()}
        self.t = 0

    def reset(self):
        self.t = 0
        if self.ens is not None:
            self.ens.reset()

    @torch.inference_mode()
    def step(self, qpos: torch.Tensor, image: torch.Tensor) -> torch.Tensor:
        """qpos [E_local, S] (raw if stats given, else already normalised), image [E_local, K, 3, H, W] in [0,1]
        -> target action [E_local, A] float64 (de-normalised if stats given)."""
        if self.stats is not None:
            qpos = (qpos - self.stats["qpos_mean"]) / self.stats["qpos_std"]
        if self.ens is not None:
            chunks = self.policy(qpos, image)                               # [E, nq, A], queried every step (:308-309)
            raw = self.ens.step(chunks, self.t)
        else:                                                               # open-loop chunk replay (:397-415)
            if self.t % self.nq == 0:
                self._chunks = self.policy(qpos, image)
            raw = self._chunks[:, self.t % self.nq].double()
        self.t += 1
        if self.stats is not None:
            raw = raw * self.stats["action_std"].double() + self.stats["action_mean"].double()
        return raw
