"""Host-side producers of the hot path's side inputs: the 2-D RoPE table (with VisionNTK extrapolation) and the
sampling script's padding helpers.

Interfaces follow FiT/preprocess/pos_embed.py:51-75 (``precompute_freqs_cis_2d``) and the helpers of
FiT/train.py:97-166 (``_pad_latent`` / ``_unpad_latent`` / ``_create_pos_embed`` / ``_create_mask``).  They are
computed once per resolution on the host in numpy -- input data, not the hot path -- and must reproduce the
reference's dtype flow exactly: with ``max_length`` the NTK-scaled theta is a float64 scalar, so the
frequencies are float64; without it they stay float32 (SURVEY.md 8c "version subtlety").
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np
import torch


def _axis_cis(axis_dim: int, coords: np.ndarray, theta: float, max_length: Optional[int]) -> np.ndarray:
    """(M, axis_dim/2, 2) of (cos, sin) for one image axis; VisionNTK scales theta by s^(axis_dim/(axis_dim-2)) with
    s = max(max(coord)/sqrt(max_length), 1)  (pos_embed.py:117-130)."""
    if max_length is not None:
        stretch = max(np.max(coords) / np.sqrt(max_length), 1.0)
        theta = theta * np.power(stretch, axis_dim / (axis_dim - 2))
    exponents = np.arange(0, axis_dim, 2, dtype=np.float32)[: axis_dim // 2] / axis_dim
    angles = np.outer(coords, 1.0 / (theta ** exponents))
    return np.stack([np.cos(angles), np.sin(angles)], axis=-1)


def precompute_freqs_cis_2d(dim: int, nh: int, nw: Optional[int] = None, theta: float = 10000.0,
                            max_length: Optional[int] = None) -> np.ndarray:
    """(nh*nw, dim) table, row = token in row-major (h, w) order.  The first dim/2 entries are the interleaved
    (cos, sin) pairs of the token's w coordinate, the last dim/2 those of its h coordinate."""
    nw = nh if nw is None else nw
    ww, hh = np.meshgrid(np.arange(nw, dtype=np.float32), np.arange(nh, dtype=np.float32))
    half = dim // 2
    table = np.concatenate([_axis_cis(half, ww.reshape(-1), theta, max_length),
                            _axis_cis(half, hh.reshape(-1), theta, max_length)], axis=1)
    return table.reshape(table.shape[0], -1)


# ---------------------------------------------------------------------------------------------- sampling-script helpers
def patchify(x: torch.Tensor, p: int) -> torch.Tensor:
    n, c, h, w = x.shape
    return x.reshape(n, c, h // p, p, w // p, p).permute(0, 2, 4, 3, 5, 1).reshape(n, (h // p) * (w // p), p * p * c)


def unpatchify(x: torch.Tensor, nh: int, nw: int, p: int, c: int) -> torch.Tensor:
    n = x.shape[0]
    return x.reshape(n, nh, nw, p, p, c).permute(0, 5, 1, 3, 2, 4).reshape(n, c, nh * p, nw * p)


def pad_latent(x: torch.Tensor, p: int, max_size: int, max_length: int) -> torch.Tensor:
    """(N,C,H,W) -> (N,C,max_size,max_size): tokens zero-padded to max_length and re-laid as a square latent;
    returned unchanged when it has more than max_length tokens (FiT/train.py:115-126)."""
    n, c = x.shape[:2]
    tok = patchify(x, p)
    if tok.shape[1] > max_length:
        return x
    buf = torch.zeros((n, max_length, p * p * c), dtype=x.dtype, device=x.device)
    buf[:, : tok.shape[1]] = tok
    return unpatchify(buf, max_size // p, max_size // p, p, c)


def unpad_latent(x: torch.Tensor, valid_t: int, h: int, w: int, p: int) -> torch.Tensor:
    """Inverse of pad_latent for the first valid_t tokens (FiT/train.py:128-134)."""
    return unpatchify(patchify(x, p)[:, :valid_t], h // p, w // p, p, x.shape[1])


def create_pos_embed(h: int, w: int, p: int, max_length: int, embed_dim: int, device=None) -> Tuple[torch.Tensor, int]:
    """(1, max(T, max_length), embed_dim) RoPE table zero-padded to max_length rows + the valid token count
    (FiT/train.py:136-156; VisionNTK always on)."""
    fill = torch.tensor(precompute_freqs_cis_2d(embed_dim, h // p, w // p, max_length=max_length), dtype=torch.float32)
    if fill.shape[0] > max_length:
        table = fill
    else:
        table = torch.zeros((max_length, embed_dim), dtype=torch.float32)
        table[: fill.shape[0]] = fill
    table = table[None]
    return (table.to(device) if device is not None else table), fill.shape[0]


def create_mask(valid_t: int, max_length: int, n: int, device=None) -> torch.Tensor:
    """(n, max(valid_t, max_length)) bool, valid prefix (FiT/train.py:158-166)."""
    if valid_t > max_length:
        m = torch.ones((valid_t,), dtype=torch.bool)
    else:
        m = torch.zeros((max_length,), dtype=torch.bool)
        m[:valid_t] = True
    m = m[None].repeat(n, 1)
    return m.to(device) if device is not None else m
