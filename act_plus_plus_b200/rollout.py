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


class DeviceRollout:
    """The rollout loop's per-step device work for E environments on ONE engine, with every pre/post-processing step of
    imitate_episodes.py fused into kernels that already run (no eager torch op per step):
      * `pre_process` (qpos z-score, :290, 370) inside the proprio projection's operand load (act_engine_set_qpos_norm),
      * uint8 HWC frames -> /255 -> CHW -> ImageNet normalise inside the stem's image pack (utils.py:133-158, policy.py:268),
      * `post_process` (action * std + mean, :294, 436) inside the ensemble / chunk-select kernel's float64 store,
      * temporal_agg (:403-411) or open-loop chunk replay (:415) with the real-robot BASE_DELAY re-timing (:311-313, 401).
    `submit()` / `step_submitted()` double-buffer the host->device copy of the NEXT observation under the current step."""

    def __init__(self, eng, num_envs: int, stats: Optional[dict] = None, temporal_agg: bool = True, base_delay: int = 0,
                 k: float = 0.01):
        c = eng.cfg
        assert num_envs <= c.batch_max, "engine workspace smaller than the number of environments"
        assert not (temporal_agg and base_delay), "the reference re-times the base only in open-loop replay (imitate_episodes.py:311-313)"
        self.eng, self.E, self.nq, self.A = eng, num_envs, c.num_queries, c.action_dim
        self.temporal_agg, self.base_delay, self.k = temporal_agg, int(base_delay), float(k)
        self.device = eng.device
        self.lib = _lib.lib()
        self.ring = torch.zeros((num_envs, self.nq, self.nq, self.A), dtype=torch.float32, device=self.device) if temporal_agg else None
        self.out = torch.empty((num_envs, self.A), dtype=torch.float64, device=self.device)
        self.t = torch.zeros(num_envs, dtype=torch.int32, device=self.device)
        self.t_host = 0
        self.chunks = None
        self.a_mean = self.a_std = None
        if stats is not None:
            f = lambda v: torch.as_tensor(v, dtype=torch.float32).to(self.device).contiguous()
            self.a_mean, self.a_std = f(stats["action_mean"]), f(stats["action_std"])
            self._qm, self._qs = f(stats["qpos_mean"]), f(stats["qpos_std"])
            check(self.lib.act_engine_set_qpos_norm(eng._h, self._qm.data_ptr(), self._qs.data_ptr(),
                                                    torch.cuda.current_stream().cuda_stream), "act_engine_set_qpos_norm")
            check(self.lib.act_engine_enable_qpos_norm(eng._h, 0), "act_engine_enable_qpos_norm")   # on only around our queries
        self._copy_stream, self._slots, self._head, self._tail = None, [], 0, 0

    def reset(self):
        self.t.zero_(); self.t_host = 0
        if self.ring is not None:
            self.ring.zero_()

    def step(self, qpos: torch.Tensor, image: torch.Tensor) -> torch.Tensor:
        """qpos [E, S] raw (with stats) or normalised, image float [E, K, 3, H, W] in [0, 1] or uint8 [E, K, H, W, 3]
        -> target action [E, A] float64 on the device (de-normalised when stats were given)."""
        st = torch.cuda.current_stream().cuda_stream
        am = self.a_mean.data_ptr() if self.a_mean is not None else None
        asd = self.a_std.data_ptr() if self.a_std is not None else None
        qf = 1 if self.temporal_agg else self.nq - self.base_delay
        if self.t_host % qf == 0:
            if self.a_mean is not None:
                check(self.lib.act_engine_enable_qpos_norm(self.eng._h, 1), "act_engine_enable_qpos_norm")
            try:
                self.chunks = self.eng.infer_graphed(qpos, image, tag=("rollout", self.a_mean is not None))        # [E, nq, A]
            finally:
                if self.a_mean is not None:
                    check(self.lib.act_engine_enable_qpos_norm(self.eng._h, 0), "act_engine_enable_qpos_norm")
        if self.temporal_agg:
            check(self.lib.act_temporal_agg_step_ex(self.ring.data_ptr(), self.chunks.data_ptr(), self.t.data_ptr(),
                                                    self.out.data_ptr(), self.E, self.nq, self.A, C.c_double(self.k), am, asd, st),
                  "act_temporal_agg_step_ex")
        else:
            check(self.lib.act_chunk_select(self.chunks.data_ptr(), self.t.data_ptr(), self.out.data_ptr(), self.E, self.nq,
                                            self.A, self.base_delay, am, asd, st), "act_chunk_select")
        self.t += 1
        self.t_host += 1
        return self.out

    # ---- host-fed loop: pinned observations in, double-buffered
    def submit(self, h_qpos: torch.Tensor, h_image: torch.Tensor):
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
            self._slots = [dict(qpos=torch.empty(h_qpos.shape, dtype=h_qpos.dtype, device=self.device),
                                image=torch.empty(h_image.shape, dtype=h_image.dtype, device=self.device),
                                ready=torch.cuda.Event(), free=None) for _ in range(2)]
        assert self._tail - self._head < 2, "two observations are already in flight: call step_submitted() first"
        s = self._slots[self._tail % 2]
        with torch.cuda.stream(self._copy_stream):
            if s["free"] is not None:
                self._copy_stream.wait_event(s["free"])
            s["qpos"].copy_(h_qpos, non_blocking=True)
            s["image"].copy_(h_image, non_blocking=True)
            s["ready"].record(self._copy_stream)
        self._tail += 1

    def step_submitted(self) -> torch.Tensor:
        assert self._head < self._tail, "nothing submitted"
        s = self._slots[self._head % 2]
        self._head += 1
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(s["ready"])
        out = self.step(s["qpos"], s["image"])
        ev = torch.cuda.Event(); ev.record(cur); s["free"] = ev
        return out


class BatchedRollout:
    """Policy-level wrapper (imitate_episodes.py:307-310, 338-339, 368-438) for E environments sharded statically across
    ranks: this rank's [lo, hi) environments run through one DeviceRollout on the policy's inference
    engine; the only host traffic per step is observations in, [E_local, A] float64 actions out."""

    def __init__(self, policy, num_envs_global: int, rank: int = 0, world: int = 1, stats: Optional[dict] = None,
                 temporal_agg: bool = True, base_delay: int = 0):
        self.policy = policy
        self.lo, self.hi = shard_range(num_envs_global, rank, world)
        self.E = self.hi - self.lo
        m = policy.model
        self.nq, self.A = m.num_queries, m.action_dim
        self.temporal_agg, self.base_delay, self.stats = temporal_agg, base_delay, stats
        self._dev = None
        self.t = 0

    def _device_rollout(self, qpos, image) -> DeviceRollout:
        eng = self.policy.model._engine_for(qpos, image, False)
        if self._dev is None or self._dev.eng is not eng:
            self._dev = DeviceRollout(eng, self.E, self.stats, self.temporal_agg, self.base_delay)
            self._dev.t.fill_(self.t); self._dev.t_host = self.t
        return self._dev

    def reset(self):
        self.t = 0
        if self._dev is not None:
            self._dev.reset()

    @torch.inference_mode()
    def step(self, qpos: torch.Tensor, image: torch.Tensor) -> torch.Tensor:
        """qpos [E_local, S] (raw if stats given, else already normalised), image [E_local, K, 3, H, W] in [0,1] (or uint8
        [E_local, K, H, W, 3]) -> target action [E_local, A] float64 (de-normalised if stats given)."""
        out = self._device_rollout(qpos, image).step(qpos, image)
        self.t += 1
        return out
