"""Data ingest for the training hot path: the contract of `EpisodicDataset.__getitem__`
(/root/reference/utils.py:60-175) and of `BatchSampler` (utils.py:239-248) without the HDF5 reader (h5py is not part of
the hot path; its output contract is this module's input contract).

Episodes live in ONE pool in their native layout -- uint8 HWC frames, raw qpos, raw actions -- either in pinned host
memory or resident in HBM (50 episodes x 400 steps x 0.92 MB = 18 GB of a B200's 180 GB).  A batch is a gather of
`batch_size` (episode, start_ts) transitions: frames stay uint8 [B, K, H, W, 3] (the engine fuses /255, the CHW permute and
the ImageNet normalisation into the stem), qpos / actions are z-scored once when an episode is added, the action window
[start_ts, start_ts + chunk) is padded with zeros exactly like utils.py:116-124 and `is_pad` marks the tail.
The reference feeds its DataLoader with 2 workers (utils.py:295); here the sampler is a few numpy calls and the gather
one indexed copy, so a data-parallel step at thousands of samples/s is not loader-bound."""
from __future__ import annotations

from typing import Dict, Iterator, List, Optional, Sequence

import numpy as np
import torch


def batch_sampler(batch_size: int, episode_len_l: Sequence[Sequence[int]], sample_weights: Optional[Sequence[float]] = None,
                  rng=None) -> Iterator[List[int]]:
    """utils.py:239-248: choose a dataset with probability `sample_weights` (normalised), then a uniform flat step index
    inside that dataset.  Draws from numpy's global RandomState in the reference's order when `rng` is None, so the same
    `np.random.seed` gives the same index stream as the reference."""
    rs = np.random if rng is None else rng
    probs = np.array(sample_weights) / np.sum(sample_weights) if sample_weights is not None else None
    bounds = np.cumsum([0] + [np.sum(episode_len) for episode_len in episode_len_l])
    while True:
        batch = []
        for _ in range(batch_size):
            d = rs.choice(len(episode_len_l), p=probs)
            batch.append(int(rs.randint(bounds[d], bounds[d + 1])))
        yield batch


class EpisodePool:
    """Episodes in native layout + the index arithmetic of EpisodicDataset (`_locate_transition`, utils.py:47-52)."""

    def __init__(self, norm_stats: Dict[str, np.ndarray], chunk_size: int, device="cpu", pin: bool = False):
        self.stats = {k: torch.as_tensor(np.asarray(v), dtype=torch.float32) for k, v in norm_stats.items()}
        self.chunk = int(chunk_size)
        self.device = torch.device(device)
        self.pin = bool(pin) and self.device.type == "cpu"
        self.frames: List[torch.Tensor] = []      # per episode uint8 [T, K, H, W, 3]
        self.qpos: List[torch.Tensor] = []        # per episode f32 [T, S], z-scored
        self.actions: List[torch.Tensor] = []     # per episode f32 [T_a, A], z-scored
        self.is_sim: List[bool] = []
        self.lens: List[int] = []
        self._cat = None

    def add_episode(self, frames_u8, qpos, actions, is_sim: bool = True):
        """frames uint8 [T, K, H, W, 3] (the HDF5 layout per camera, stacked on K), qpos [T, S], actions [T, A] raw."""
        f = torch.as_tensor(frames_u8)
        assert f.dtype == torch.uint8 and f.dim() == 5 and f.shape[-1] == 3
        q = (torch.as_tensor(qpos, dtype=torch.float32) - self.stats["qpos_mean"]) / self.stats["qpos_std"]       # utils.py:166
        a = (torch.as_tensor(actions, dtype=torch.float32) - self.stats["action_mean"]) / self.stats["action_std"]  # utils.py:164
        put = (lambda t: t.contiguous().pin_memory()) if self.pin else (lambda t: t.contiguous().to(self.device))
        self.frames.append(put(f)); self.qpos.append(put(q)); self.actions.append(put(a))
        self.is_sim.append(bool(is_sim)); self.lens.append(int(f.shape[0]))
        self._cat = None

    def __len__(self):
        return int(sum(self.lens))

    def _build(self):
        cum = np.cumsum([0] + self.lens)
        self._cum = cum
        self._act_cum = np.cumsum([0] + [int(a.shape[0]) for a in self.actions])      # host-side: no device sync per item
        self._cat = dict(frames=torch.cat(self.frames), qpos=torch.cat(self.qpos), actions=torch.cat(self.actions))
        if self.pin:
            self._cat = {k: (v.pin_memory() if v.device.type == "cpu" else v) for k, v in self._cat.items()}

    def locate(self, index: int):
        """utils.py:47-52: flat step index -> (episode, start_ts)."""
        if self._cat is None:
            self._build()
        e = int(np.argmax(self._cum > index)) - 1
        return e, int(index - self._cum[e])

    def gather(self, indices: Sequence[int], out: Optional[Dict[str, torch.Tensor]] = None) -> Dict[str, torch.Tensor]:
        """One batch in the layout the engine ingests: image uint8 [B, K, H, W, 3], qpos f32 [B, S], actions f32
        [B, chunk, A] (rows past the episode end hold the z-scored zero padding of utils.py:116-117, 164), is_pad bool [B, chunk].  `out` (optional) receives the batch in
        place -- e.g. the pinned staging buffers an InputPrefetcher copies from."""
        if self._cat is None:
            self._build()
        c, dev = self._cat, self._cat["frames"].device
        loc = [self.locate(i) for i in indices]
        flat = torch.as_tensor([self._cum[e] + ts for e, ts in loc], device=dev)
        # first action row of the window and number of valid rows (utils.py:108-114: real data starts one step earlier)
        first = [ts if self.is_sim[e] else max(0, ts - 1) for e, ts in loc]
        alen = [int(self.actions[e].shape[0]) - f for (e, _), f in zip(loc, first)]
        a0 = torch.as_tensor([int(self._act_cum[e]) + f for (e, _), f in zip(loc, first)], device=dev)
        steps = torch.arange(self.chunk, device=dev)
        valid = steps[None, :] < torch.as_tensor(alen, device=dev)[:, None]
        rows = (a0[:, None] + steps[None, :]).clamp_(max=c["actions"].shape[0] - 1)
        # the reference pads the RAW window with zeros and z-scores afterwards (utils.py:116-117, 164): a padded row holds
        # (0 - mean) / std, not 0 (it is masked everywhere downstream, but the contract is kept literally)
        pad_row = ((0.0 - self.stats["action_mean"]) / self.stats["action_std"]).to(dev)
        actions = torch.where(valid[..., None], c["actions"][rows], pad_row)
        batch = dict(image=c["frames"].index_select(0, flat), qpos=c["qpos"].index_select(0, flat), actions=actions,
                     is_pad=~valid)
        if out is None:
            return batch
        for k, v in batch.items():
            out[k].copy_(v)
        return out
