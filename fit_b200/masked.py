"""masked_FiT batch assembly on the device.

``ImageNetLatentIterator.collate`` (masked_FiT/preprocess/iterators.py:157-169) draws ``number_of_tokens`` from {32, 64, 96, 128, 192, 256}
and calls ``subsample`` (iterators.py:142-155) per sample: samples with more tokens keep a random subset in permuted order together
with their RoPE rows, the others are zero-padded (mask False).  Here the random draws stay on the host in the reference's order
(``random.choice`` by the caller, one ``torch.randperm(n)`` per over-long sample, in batch order) and the gather of latent tokens, RoPE
rows and mask bytes of the whole batch is one kernel (``fitb200_subsample_tokens``).  No CPU fallback."""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch

from . import ops

TOKEN_CHOICES = (32, 64, 96, 128, 192, 256)   # iterators.py:158


def subsample_batch(latents: Sequence[torch.Tensor], poss: Sequence[torch.Tensor], masks: Sequence[torch.Tensor], number_of_tokens: int,
                    device, generator: Optional[torch.Generator] = None) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """latents[i]: [T_i, P] fp32, poss[i]: [T_i, d] fp32, masks[i]: [T_i] bool (host tensors, as the dataset yields them).
    Returns device tensors (latent [N, T', P], pos [N, T', d], mask [N, T'])."""
    N = len(latents)
    Tsrc = max(int(x.shape[0]) for x in latents)
    P, d = int(latents[0].shape[1]), int(poss[0].shape[1])
    lat = torch.zeros(N, Tsrc, P)
    pos = torch.zeros(N, Tsrc, d)
    msk = torch.zeros(N, Tsrc, dtype=torch.bool)
    perm = torch.zeros(N, Tsrc, dtype=torch.int64)
    n_tok = torch.zeros(N, dtype=torch.int32)
    for i, (l, p, m) in enumerate(zip(latents, poss, masks)):
        n = int(l.shape[0])
        lat[i, :n], pos[i, :n], msk[i, :n], n_tok[i] = l, p, m, n
        if n > number_of_tokens:   # iterators.py:143-144: the permutation is drawn only for over-long samples
            perm[i, :n] = torch.randperm(n, generator=generator)
    dev = torch.device(device)
    up = lambda t: t.pin_memory().to(dev, non_blocking=True)  # noqa: E731
    return ops.subsample_tokens(up(lat), up(pos), up(msk), up(perm), up(n_tok), number_of_tokens)
