"""Seeded synthetic inputs in the reference iterators' output contract (TEST INFRASTRUCTURE).

FiT/preprocess/iterators.py:141-166: latent tokens zero-padded to T, RoPE rows zero-padded,
mask = valid prefix.  masked_FiT/preprocess/iterators.py:142-169: random token subset in
permuted order with their RoPE rows (``masked_batch``).
Used by oracle/make_goldens.py (with the reference's table generator) and by tests/ (with the
oracle's); the input checksums stored in the fixtures prove both produce identical inputs.
"""
from __future__ import annotations

import torch

from . import fit_oracle as fo
from .pos_embed_oracle import precompute_freqs_cis_2d


def mixed_batch(cfg, shapes, T, seed, rope_dim, precompute=precompute_freqs_cis_2d):
    g = torch.Generator().manual_seed(seed)
    N = len(shapes)
    p = cfg["patch_size"]
    x = torch.zeros(N, T, p * p * cfg["in_channels"])
    pos = torch.zeros(N, T, rope_dim)
    mask = torch.zeros(N, T, dtype=torch.bool)
    for i, (h, w) in enumerate(shapes):
        lat = torch.randn(1, cfg["in_channels"], h, w, generator=g)
        tok = fo.patchify(lat, p)[0]
        L = tok.shape[0]
        x[i, :L] = tok
        pos[i, :L] = torch.tensor(precompute(rope_dim, h // p, w // p, max_length=T), dtype=torch.float32)
        mask[i, :L] = True
    t = torch.randint(0, 1000, (N,), generator=g)
    y = torch.randint(0, cfg["num_classes"], (N,), generator=g)
    target = torch.randn(N, T, p * p * cfg["in_channels"], generator=g)
    return x, t, y, pos, mask, target


def eval_batch(cfg, h, w, T_ntk, seed=13, B=2, t_value=333, precompute=precompute_freqs_cis_2d):
    """CFG-style eval inputs: (2B,C,h,w) duplicated latents, labels | null class."""
    d = cfg["hidden_size"] // cfg["num_heads"]
    p = cfg["patch_size"]
    g = torch.Generator().manual_seed(seed)
    z = torch.randn(B, cfg["in_channels"], h, w, generator=g)
    zz = torch.cat([z, z], 0)
    yy = torch.cat([torch.randint(0, cfg["num_classes"], (B,), generator=g),
                    torch.full((B,), cfg["num_classes"])])
    Te = (h // p) * (w // p)
    pos = torch.tensor(precompute(d, h // p, w // p, max_length=T_ntk), dtype=torch.float32)[None]
    mask = torch.ones(2 * B, Te, dtype=torch.bool)
    tt = torch.full((2 * B,), t_value, dtype=torch.long)
    return zz, tt, yy, pos, mask


def masked_batch(cfg, shapes, n_tokens, seed, rope_dim, T_ntk=256):
    """masked_FiT collate: every sample reduced to a random permutation of its tokens truncated
    to n_tokens, or zero-padded up to it (masked_FiT/preprocess/iterators.py:142-169)."""
    g = torch.Generator().manual_seed(seed)
    p = cfg["patch_size"]
    N = len(shapes)
    x = torch.zeros(N, n_tokens, p * p * cfg["in_channels"])
    pos = torch.zeros(N, n_tokens, rope_dim)
    mask = torch.zeros(N, n_tokens, dtype=torch.bool)
    for i, (h, w) in enumerate(shapes):
        tok = fo.patchify(torch.randn(1, cfg["in_channels"], h, w, generator=g), p)[0]
        tab = torch.tensor(precompute_freqs_cis_2d(rope_dim, h // p, w // p, max_length=T_ntk), dtype=torch.float32)
        L = tok.shape[0]
        if L > n_tokens:
            idx = torch.randperm(L, generator=g)[:n_tokens]
            x[i], pos[i], mask[i] = tok[idx], tab[idx], True
        else:
            x[i, :L], pos[i, :L], mask[i, :L] = tok, tab, True
    t = torch.randint(0, 1000, (N,), generator=g)
    y = torch.randint(0, cfg["num_classes"], (N,), generator=g)
    target = torch.randn(N, n_tokens, p * p * cfg["in_channels"], generator=g)
    return x, t, y, pos, mask, target


# ---- masked_FiT fixture (tests/golden/masked_fit_tiny.npz, written by oracle/make_goldens.py::gen_masked from the reference's own
# ImageNetLatentIterator._patchify / subsample / collate and training step)
MASKED_SHAPES = [(32, 32), (32, 32), (20, 40), (8, 16)]   # 256, 256, 200 tokens (sub-sampled to 128) and 32 tokens (padded)
MASKED_CFG = dict(patch_size=2, in_channels=4, hidden_size=144, depth=2, num_heads=2, mlp_ratio=4.0,
                  class_dropout_prob=0.1, num_classes=10, learn_sigma=False)   # head_dim 72, like FiT-XL/2


def masked_inputs(seed=31):
    """Seeded dataset items in the form masked_FiT's __getitem__ hands to collate (iterators.py:171-186): raw latents + labels."""
    g = torch.Generator().manual_seed(seed)
    lats = [torch.randn(4, h, w, generator=g).numpy() for (h, w) in MASKED_SHAPES]
    labels = [int(v) for v in torch.randint(0, 10, (len(MASKED_SHAPES),), generator=g)]
    return lats, labels


def masked_items(seed=31):
    """The same items after the iterator's _patchify (iterators.py:128-140): token rows, RoPE rows (table built WITHOUT max_length,
    i.e. fp32-precision frequencies, as the iterator does), all-ones masks.  Returns (latents, poss, masks, labels) as host tensors."""
    lats, labels = masked_inputs(seed)
    toks, poss, masks = [], [], []
    for lat in lats:
        tok = fo.patchify(torch.tensor(lat)[None], 2)[0]
        h, w = lat.shape[1] // 2, lat.shape[2] // 2
        toks.append(tok.contiguous())
        poss.append(torch.tensor(precompute_freqs_cis_2d(72, h, w).astype("float32")))
        masks.append(torch.ones(tok.shape[0], dtype=torch.bool))
    return toks, poss, masks, labels


def masked_step_noise(shape, seed=17):
    """t and noise of the masked training-step fixture (drawn in the order gen_masked draws them)."""
    g = torch.Generator().manual_seed(seed)
    t = torch.randint(0, 1000, (shape[0],), generator=g)
    noise = torch.randn(shape, generator=g)
    return t, noise
