"""Constant tables of the ACT model that are cheap enough to build on the host."""
import numpy as np
import torch


def sinusoid_table(n_position: int, d_hid: int) -> torch.Tensor:
    """detr_vae.py:25-33: float64 angle table, sin on even / cos on odd columns, cast to float32 -> [1, n, d]."""
    pos = np.arange(n_position, dtype=np.float64)[:, None]
    j = np.arange(d_hid)[None, :]
    table = pos / np.power(10000, 2 * (j // 2) / d_hid)
    table[:, 0::2] = np.sin(table[:, 0::2])
    table[:, 1::2] = np.cos(table[:, 1::2])
    return torch.from_numpy(table.astype(np.float32)).unsqueeze(0)
