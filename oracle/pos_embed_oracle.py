"""numpy restatement of the 2-D RoPE table generator (TEST INFRASTRUCTURE).

Follows FiT/preprocess/pos_embed.py:51-75 (precompute_freqs_cis_2d) and :108-130
(_precompute_freqs_cis_2d_from_grid / _1d_from_grid, incl. VisionNTK).
Pinned against the reference by tests/golden/rope_tables.npz.
"""
from __future__ import annotations

from typing import Optional

import numpy as np


def _axis_table(dim: int, pos: np.ndarray, theta: float, max_length: Optional[int]) -> np.ndarray:
    """One axis: (M, dim/2, 2) of (cos, sin) -- pos_embed.py:117-130.

    VisionNTK (pos_embed.py:120-123): s = max(max(pos)/sqrt(max_length), 1),
    theta <- theta * s**(dim/(dim-2)).  `dim` here is the per-axis dim (= head_dim/2).
    Note the dtype flow of the reference: with max_length the scaled theta is a numpy
    float64 scalar, so the frequencies are float64; without it they are float32."""
    if max_length is not None:
        s = max(np.max(pos) / np.sqrt(max_length), 1.0)
        theta = theta * np.power(s, dim / (dim - 2))
    expo = np.arange(0, dim, 2, dtype=np.float32)[: dim // 2] / dim
    inv = 1.0 / (theta ** expo)
    ang = np.outer(pos, inv)
    return np.stack([np.cos(ang), np.sin(ang)], axis=-1)


def precompute_freqs_cis_2d(dim: int, nh: int, nw: Optional[int] = None, theta: float = 10000.0,
                            max_length: Optional[int] = None) -> np.ndarray:
    """(nh*nw, dim) table; row = token in row-major (h, w) order; first dim/2 entries are the
    interleaved (cos, sin) pairs of the w coordinate, last dim/2 those of the h coordinate
    -- pos_embed.py:51-75, 108-114."""
    nw = nh if nw is None else nw
    gw, gh = np.meshgrid(np.arange(nw, dtype=np.float32), np.arange(nh, dtype=np.float32))
    tw = _axis_table(dim // 2, gw.reshape(-1), theta, max_length)
    th = _axis_table(dim // 2, gh.reshape(-1), theta, max_length)
    out = np.concatenate([tw, th], axis=1)
    return out.reshape(out.shape[0], -1)
