"""Timestep respacing -- mirror of FiT/diffusion/respace.py (space_timesteps :12-62, SpacedDiffusion :65-115,
_WrappedModel :117-129).  The loop-index -> original-timestep map lives on the device and is applied by
fitb200_gather_i64 (no per-step host-to-device copy, respace.py:125)."""
from __future__ import annotations

import numpy as np
import torch

from .. import ops
from .gaussian_diffusion import GaussianDiffusion


def space_timesteps(num_timesteps, section_counts):
    """Set of original timesteps to keep.  ``section_counts``: "250", "10,10,5", [..] or "ddimN" (fixed stride)."""
    if isinstance(section_counts, str):
        if section_counts.startswith("ddim"):
            want = int(section_counts[len("ddim"):])
            for stride in range(1, num_timesteps):
                if len(range(0, num_timesteps, stride)) == want:
                    return set(range(0, num_timesteps, stride))
            raise ValueError(f"cannot create exactly {num_timesteps} steps with an integer stride")
        section_counts = [int(x) for x in section_counts.split(",")]
    n_sec = len(section_counts)
    base, extra = divmod(num_timesteps, n_sec)
    kept, offset = [], 0
    for k, count in enumerate(section_counts):
        size = base + (1 if k < extra else 0)
        if size < count:
            raise ValueError(f"cannot divide section of {size} steps into {count}")
        step = 1.0 if count <= 1 else (size - 1) / (count - 1)
        pos = 0.0  # accumulated (not multiplied) so that ties round exactly as in the reference
        for _ in range(count):
            kept.append(offset + round(pos))
            pos += step
        offset += size
    return set(kept)


class SpacedDiffusion(GaussianDiffusion):
    """A diffusion process that skips steps of a base process: betas re-derived from the base alphas_cumprod at the
    kept steps, and the model always sees ORIGINAL timestep values."""

    def __init__(self, use_timesteps, **kwargs):
        self.use_timesteps = set(use_timesteps)
        self.original_num_steps = len(kwargs["betas"])
        base_ac = np.cumprod(1.0 - np.asarray(kwargs["betas"], dtype=np.float64))
        self.timestep_map = sorted(i for i in self.use_timesteps if 0 <= i < self.original_num_steps)
        prev, new_betas = 1.0, []
        for i in self.timestep_map:
            new_betas.append(1 - base_ac[i] / prev)
            prev = base_ac[i]
        kwargs["betas"] = np.array(new_betas)
        super().__init__(**kwargs)
        self._map_dev = {}

    def _map_tensor(self, device):
        m = self._map_dev.get(str(device))
        if m is None:
            m = torch.tensor(self.timestep_map, dtype=torch.int64, device=device)
            self._map_dev[str(device)] = m
        return m

    def _model_timesteps(self, t):
        t = t.long()
        if not t.is_cuda:
            raise RuntimeError("fit_b200: timesteps must be CUDA tensors (no CPU fallback)")
        return ops.gather_i64(self._map_tensor(t.device), t.contiguous(), torch.empty_like(t))

    def _wrap_model(self, model):
        if isinstance(model, _WrappedModel):
            return model
        return _WrappedModel(model, self, self.original_num_steps)

    def _scale_timesteps(self, t):
        return t


class _WrappedModel:
    def __init__(self, model, diffusion, original_num_steps):
        self.model = model
        self.diffusion = diffusion
        self.original_num_steps = original_num_steps

    def __call__(self, x, ts, **kwargs):
        return self.model(x, self.diffusion._model_timesteps(ts), **kwargs)
