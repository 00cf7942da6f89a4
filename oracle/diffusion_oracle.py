"""CPU restatement of the diffusion step arithmetic (TEST INFRASTRUCTURE).

float64 numpy coefficient tables exactly as FiT/diffusion/gaussian_diffusion.py:165-213 and
FiT/diffusion/respace.py:12-92 build them; the per-step update in fp32 torch as
p_mean_variance (:266-344), p_sample (:388-429), ddim_sample (:525-572), q_sample (:227-242).
Configuration fixed to what create_diffusion() produces by default
(FiT/diffusion/__init__.py:10-46): linear betas, EPSILON, FIXED_LARGE, MSE.
Pinned against the reference by tests/golden/diffusion_tables.npz and diffusion_steps.npz.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional

import numpy as np
import torch

Tensor = torch.Tensor


def linear_betas(n: int) -> np.ndarray:
    """get_named_beta_schedule("linear", n) -- gaussian_diffusion.py:110-127, 91-92."""
    scale = 1000 / n
    return np.linspace(scale * 0.0001, scale * 0.02, n, dtype=np.float64)


def space_timesteps(num_timesteps: int, section_counts) -> List[int]:
    """Sorted kept timesteps -- respace.py:12-62 (incl. the "ddimN" form)."""
    if isinstance(section_counts, str):
        if section_counts.startswith("ddim"):
            want = int(section_counts[4:])
            for stride in range(1, num_timesteps):
                if len(range(0, num_timesteps, stride)) == want:
                    return sorted(range(0, num_timesteps, stride))
            raise ValueError(f"cannot create exactly {num_timesteps} steps with an integer stride")
        section_counts = [int(v) for v in section_counts.split(",")]
    per, extra = divmod(num_timesteps, len(section_counts))
    start, keep = 0, set()
    for i, cnt in enumerate(section_counts):
        size = per + (1 if i < extra else 0)
        if size < cnt:
            raise ValueError(f"cannot divide section of {size} steps into {cnt}")
        stride = 1 if cnt <= 1 else (size - 1) / (cnt - 1)
        cur = 0.0
        for _ in range(cnt):
            keep.add(start + round(cur))
            cur += stride
        start += size
    return sorted(keep)


class Tables:
    """All float64 tables of a (respaced) diffusion -- gaussian_diffusion.py:179-213."""

    def __init__(self, betas: np.ndarray):
        b = np.asarray(betas, dtype=np.float64)
        self.betas = b
        self.num_timesteps = len(b)
        a = 1.0 - b
        self.alphas_cumprod = np.cumprod(a)
        self.alphas_cumprod_prev = np.append(1.0, self.alphas_cumprod[:-1])
        self.sqrt_alphas_cumprod = np.sqrt(self.alphas_cumprod)
        self.sqrt_one_minus_alphas_cumprod = np.sqrt(1.0 - self.alphas_cumprod)
        self.sqrt_recip_alphas_cumprod = np.sqrt(1.0 / self.alphas_cumprod)
        self.sqrt_recipm1_alphas_cumprod = np.sqrt(1.0 / self.alphas_cumprod - 1)
        self.posterior_variance = b * (1.0 - self.alphas_cumprod_prev) / (1.0 - self.alphas_cumprod)
        self.posterior_log_variance_clipped = np.log(np.append(self.posterior_variance[1], self.posterior_variance[1:]))
        self.posterior_mean_coef1 = b * np.sqrt(self.alphas_cumprod_prev) / (1.0 - self.alphas_cumprod)
        self.posterior_mean_coef2 = (1.0 - self.alphas_cumprod_prev) * np.sqrt(a) / (1.0 - self.alphas_cumprod)
        # FIXED_LARGE variance -- gaussian_diffusion.py:307-314
        self.fixed_large_variance = np.append(self.posterior_variance[1], b[1:])
        self.fixed_large_log_variance = np.log(self.fixed_large_variance)


def create(timestep_respacing, diffusion_steps: int = 1000):
    """(Tables, timestep_map) of create_diffusion(timestep_respacing) -- __init__.py:10-46,
    respace.py:73-87 (betas re-derived from alphas_cumprod at the kept steps)."""
    base = Tables(linear_betas(diffusion_steps))
    if timestep_respacing is None or timestep_respacing == "":
        timestep_respacing = [diffusion_steps]
    keep = space_timesteps(diffusion_steps, timestep_respacing)
    last, new_betas = 1.0, []
    for i in keep:
        new_betas.append(1 - base.alphas_cumprod[i] / last)
        last = base.alphas_cumprod[i]
    return Tables(np.array(new_betas)), keep


def _gather(arr: np.ndarray, t: Tensor, ndim: int) -> Tensor:
    """_extract_into_tensor: float64 table -> .float() per-row scalar -- gaussian_diffusion.py:882-894."""
    v = torch.from_numpy(arr).to(t.device)[t.long()].float()   # the reference copies the table to the timestep's device per call
    return v.reshape(-1, *([1] * (ndim - 1)))


def p_mean_variance(tb: Tables, x: Tensor, t: Tensor, eps: Tensor) -> Dict[str, Tensor]:
    """EPSILON / FIXED_LARGE / clip_denoised=False branch -- gaussian_diffusion.py:266-344."""
    n = x.dim()
    x0 = _gather(tb.sqrt_recip_alphas_cumprod, t, n) * x - _gather(tb.sqrt_recipm1_alphas_cumprod, t, n) * eps
    mean = _gather(tb.posterior_mean_coef1, t, n) * x0 + _gather(tb.posterior_mean_coef2, t, n) * x
    zeros = torch.zeros_like(x)
    return {"mean": mean, "pred_xstart": x0,
            "variance": _gather(tb.fixed_large_variance, t, n) + zeros,
            "log_variance": _gather(tb.fixed_large_log_variance, t, n) + zeros}


def p_sample(tb: Tables, x: Tensor, t: Tensor, eps: Tensor, noise: Tensor) -> Dict[str, Tensor]:
    """mean + (t != 0) exp(0.5 logvar) noise -- gaussian_diffusion.py:388-429."""
    out = p_mean_variance(tb, x, t, eps)
    nz = (t != 0).float().reshape(-1, *([1] * (x.dim() - 1)))
    return {"sample": out["mean"] + nz * torch.exp(0.5 * out["log_variance"]) * noise,
            "pred_xstart": out["pred_xstart"]}


def ddim_sample(tb: Tables, x: Tensor, t: Tensor, eps_model: Tensor, noise: Tensor, eta: float = 0.0):
    """DDIM update, eps re-derived from pred_xstart -- gaussian_diffusion.py:525-572, 353-356."""
    n = x.dim()
    out = p_mean_variance(tb, x, t, eps_model)
    x0 = out["pred_xstart"]
    eps = (_gather(tb.sqrt_recip_alphas_cumprod, t, n) * x - x0) / _gather(tb.sqrt_recipm1_alphas_cumprod, t, n)
    ab = _gather(tb.alphas_cumprod, t, n)
    ab_prev = _gather(tb.alphas_cumprod_prev, t, n)
    sigma = eta * torch.sqrt((1 - ab_prev) / (1 - ab)) * torch.sqrt(1 - ab / ab_prev)
    mean = x0 * torch.sqrt(ab_prev) + torch.sqrt(1 - ab_prev - sigma ** 2) * eps
    nz = (t != 0).float().reshape(-1, *([1] * (n - 1)))
    return {"sample": mean + nz * sigma * noise, "pred_xstart": x0}


def q_sample(tb: Tables, x_start: Tensor, t: Tensor, noise: Tensor) -> Tensor:
    """sqrt(abar_t) x0 + sqrt(1-abar_t) noise -- gaussian_diffusion.py:227-242.
    Also the in-repo stand-in for diffusers' DDIMScheduler.add_noise (FiT/train.py:53; parity
    unpinned for that third-party call)."""
    n = x_start.dim()
    return _gather(tb.sqrt_alphas_cumprod, t, n) * x_start + _gather(tb.sqrt_one_minus_alphas_cumprod, t, n) * noise


def sample_loop(tb: Tables, timestep_map: List[int], model: Callable, x: Tensor,
                noise_fn: Callable[[Tensor], Tensor], mode: str = "ddpm", eta: float = 0.0,
                model_kwargs: Optional[dict] = None, trace: Optional[list] = None) -> Tensor:
    """p_sample_loop / ddim_sample_loop with the _WrappedModel timestep mapping
    -- gaussian_diffusion.py:431-523, 612-692; respace.py:117-129.
    ``noise_fn(x)`` is called once per step (incl. t=0), like th.randn_like in the reference."""
    model_kwargs = model_kwargs or {}
    tmap = torch.tensor(timestep_map, dtype=torch.long)
    for i in reversed(range(tb.num_timesteps)):
        t = torch.full((x.shape[0],), i, dtype=torch.long)
        with torch.no_grad():
            eps = model(x, tmap[t], **model_kwargs)
            # reference order: p_mean_variance (model call) first, then randn_like
            noise = noise_fn(x)
            out = p_sample(tb, x, t, eps, noise) if mode == "ddpm" else ddim_sample(tb, x, t, eps, noise, eta)
        x = out["sample"]
        if trace is not None:
            trace.append(x)
    return x
