"""Mirror of the reference's ``diffusion.gaussian_diffusion`` object API on the fused CUDA step kernel.

Interface follows FiT/diffusion/gaussian_diffusion.py (GaussianDiffusion: q_sample :227, q_posterior_mean_variance
:244, p_mean_variance :266, p_sample :388, p_sample_loop(_progressive) :431/:476, ddim_sample :525,
ddim_sample_loop(_progressive) :612/:645, training_losses :727) with the same argument names and returned
dict keys.  The coefficient tables are float64 numpy (input data, computed once on the host); every per-step
tensor operation is ONE launch of fitb200_cfg_ddpm_step / fitb200_cfg_ddim_step / fitb200_q_sample --
the reference's ~10 elementwise kernels and 6 host-to-device `_extract_into_tensor` copies per step
(:882-894) are gone.  Model calls use the reference convention ``model(x, t, **model_kwargs)``.

Unsupported on purpose (not on the configured path): learned variances
(learn_sigma is disabled in the reference, fit.py:360), the KL / VLB loss family, cond_fn / denoised_fn hooks.
"""
from __future__ import annotations

import enum
import math
from typing import Dict, Optional

import numpy as np
import torch

from .. import ops

Tensor = torch.Tensor


class ModelMeanType(enum.Enum):
    PREVIOUS_X = enum.auto()
    START_X = enum.auto()
    EPSILON = enum.auto()


class ModelVarType(enum.Enum):
    LEARNED = enum.auto()
    FIXED_SMALL = enum.auto()
    FIXED_LARGE = enum.auto()
    LEARNED_RANGE = enum.auto()


class LossType(enum.Enum):
    MSE = enum.auto()
    RESCALED_MSE = enum.auto()
    KL = enum.auto()
    RESCALED_KL = enum.auto()

    def is_vb(self):
        return self in (LossType.KL, LossType.RESCALED_KL)


def get_named_beta_schedule(schedule_name: str, num_diffusion_timesteps: int) -> np.ndarray:
    """"linear" (scaled to the step count) or "squaredcos_cap_v2" -- gaussian_diffusion.py:110-153."""
    n = num_diffusion_timesteps
    if schedule_name == "linear":
        k = 1000 / n
        return np.linspace(k * 0.0001, k * 0.02, n, dtype=np.float64)
    if schedule_name == "squaredcos_cap_v2":
        def abar(u):
            return math.cos((u + 0.008) / 1.008 * math.pi / 2) ** 2
        return np.array([min(1 - abar((i + 1) / n) / abar(i / n), 0.999) for i in range(n)], dtype=np.float64)
    raise NotImplementedError(f"unknown beta schedule: {schedule_name}")


class GaussianDiffusion:
    def __init__(self, *, betas, model_mean_type, model_var_type, loss_type):
        if model_var_type not in (ModelVarType.FIXED_LARGE, ModelVarType.FIXED_SMALL):
            raise NotImplementedError("fit_b200: learned variances are disabled in the reference (learn_sigma, fit.py:360,372)")
        if model_mean_type not in (ModelMeanType.EPSILON, ModelMeanType.START_X):
            raise NotImplementedError("fit_b200: model_mean_type must be EPSILON or START_X")
        self.model_mean_type = model_mean_type
        self.model_var_type = model_var_type
        self.loss_type = loss_type
        b = np.asarray(betas, dtype=np.float64)
        assert b.ndim == 1 and (b > 0).all() and (b <= 1).all()
        self.betas = b
        self.num_timesteps = int(b.shape[0])
        ac = np.cumprod(1.0 - b)
        self.alphas_cumprod = ac
        self.alphas_cumprod_prev = np.concatenate([[1.0], ac[:-1]])
        self.alphas_cumprod_next = np.concatenate([ac[1:], [0.0]])
        self.sqrt_alphas_cumprod = np.sqrt(ac)
        self.sqrt_one_minus_alphas_cumprod = np.sqrt(1.0 - ac)
        self.log_one_minus_alphas_cumprod = np.log(1.0 - ac)
        self.sqrt_recip_alphas_cumprod = np.sqrt(1.0 / ac)
        self.sqrt_recipm1_alphas_cumprod = np.sqrt(1.0 / ac - 1)
        pv = b * (1.0 - self.alphas_cumprod_prev) / (1.0 - ac)
        self.posterior_variance = pv
        self.posterior_log_variance_clipped = np.log(np.concatenate([pv[1:2], pv[1:]])) if len(pv) > 1 else np.array([])
        self.posterior_mean_coef1 = b * np.sqrt(self.alphas_cumprod_prev) / (1.0 - ac)
        self.posterior_mean_coef2 = (1.0 - self.alphas_cumprod_prev) * np.sqrt(1.0 - b) / (1.0 - ac)
        self._dev: Dict[str, Dict[str, Tensor]] = {}

    # ------------------------------------------------------------------ device tables
    def _model_variance_tables(self):
        """(variance, log_variance) per timestep for the fixed variance types -- gaussian_diffusion.py:307-320."""
        if self.model_var_type == ModelVarType.FIXED_LARGE:
            var = np.concatenate([self.posterior_variance[1:2], self.betas[1:]])
            return var, np.log(var)
        return self.posterior_variance, self.posterior_log_variance_clipped

    def _tables(self, device) -> Dict[str, Tensor]:
        key = str(device)
        tb = self._dev.get(key)
        if tb is None:
            var, logvar = self._model_variance_tables()
            # each float64 entry cast to fp32 exactly as _extract_into_tensor does (.float(), :891)
            coef = np.stack([self.sqrt_recip_alphas_cumprod, self.sqrt_recipm1_alphas_cumprod, self.posterior_mean_coef1,
                             self.posterior_mean_coef2, logvar, self.alphas_cumprod, self.alphas_cumprod_prev, var], axis=1)
            f = lambda a: torch.from_numpy(np.ascontiguousarray(a)).float().to(device)  # noqa: E731
            tb = {"coef": f(coef), "sqrt_ac": f(self.sqrt_alphas_cumprod), "sqrt_1m_ac": f(self.sqrt_one_minus_alphas_cumprod),
                  "q_coef": f(np.stack([np.zeros_like(var)] * 2 + [self.posterior_mean_coef1, self.posterior_mean_coef2] +
                                       [np.zeros_like(var)] * 4, axis=1))}
            self._dev[key] = tb
        return tb

    def _flags(self, clip_denoised: bool) -> int:
        f = ops.STEP_CLIP_DENOISED if clip_denoised else 0
        if self.model_mean_type == ModelMeanType.START_X:
            f |= ops.STEP_PREDICT_XSTART
        return f

    @staticmethod
    def _per_row(table_col: Tensor, t: Tensor, like: Tensor) -> Tensor:
        return table_col[t].reshape(-1, *([1] * (like.dim() - 1))).expand(like.shape)

    # ------------------------------------------------------------------ forward process
    def q_mean_variance(self, x_start, t):
        tb = self._tables(x_start.device)
        zeros = torch.zeros_like(x_start)
        mean = ops.q_sample(x_start.float().contiguous(), zeros, t.long(), tb["sqrt_ac"], tb["sqrt_1m_ac"])
        one_m = torch.from_numpy(1.0 - self.alphas_cumprod).float().to(x_start.device)
        log_one_m = torch.from_numpy(self.log_one_minus_alphas_cumprod).float().to(x_start.device)
        return mean, self._per_row(one_m, t, x_start), self._per_row(log_one_m, t, x_start)

    def q_sample(self, x_start, t, noise=None):
        """sqrt(abar_t) x0 + sqrt(1 - abar_t) noise -- gaussian_diffusion.py:227-242."""
        if noise is None:
            noise = torch.randn_like(x_start)
        assert noise.shape == x_start.shape
        tb = self._tables(x_start.device)
        return ops.q_sample(x_start.float().contiguous(), noise.float().contiguous(), t.long(), tb["sqrt_ac"], tb["sqrt_1m_ac"])

    def q_posterior_mean_variance(self, x_start, x_t, t):
        """q(x_{t-1} | x_t, x_0) -- gaussian_diffusion.py:244-264."""
        assert x_start.shape == x_t.shape
        tb = self._tables(x_t.device)
        mean = torch.empty_like(x_t, dtype=torch.float32)
        ops.cfg_step(x_t.float().contiguous(), x_start.float().contiguous(), None, t.long(), tb["q_coef"],
                     flags=ops.STEP_PREDICT_XSTART, mean=mean, want_sample=False)
        pv = torch.from_numpy(self.posterior_variance).float().to(x_t.device)
        plv = torch.from_numpy(self.posterior_log_variance_clipped).float().to(x_t.device)
        return mean, self._per_row(pv, t, x_t), self._per_row(plv, t, x_t)

    # ------------------------------------------------------------------ reverse process
    def _check_hooks(self, denoised_fn, cond_fn=None):
        if denoised_fn is not None or cond_fn is not None:
            raise NotImplementedError("fit_b200: denoised_fn / cond_fn hooks are not on the reference's sampling path")

    def p_mean_variance(self, model, x, t, clip_denoised=True, denoised_fn=None, model_kwargs=None):
        """gaussian_diffusion.py:266-344: dict(mean, variance, log_variance, pred_xstart)."""
        self._check_hooks(denoised_fn)
        model_kwargs = model_kwargs or {}
        model_output = model(x, self._model_timesteps(t), **model_kwargs)
        assert model_output.shape == x.shape
        tb = self._tables(x.device)
        mean = torch.empty_like(x, dtype=torch.float32)
        x0 = torch.empty_like(x, dtype=torch.float32)
        t = t.long()
        ops.cfg_step(x.float().contiguous(), model_output.float().contiguous(), None, t, tb["coef"],
                     flags=self._flags(clip_denoised), pred_xstart=x0, mean=mean, want_sample=False)
        return {"mean": mean, "variance": self._per_row(tb["coef"][:, 7], t, x),
                "log_variance": self._per_row(tb["coef"][:, 4], t, x), "pred_xstart": x0}

    def _model_timesteps(self, t):
        return t

    def p_sample(self, model, x, t, clip_denoised=True, denoised_fn=None, cond_fn=None, model_kwargs=None, noise=None):
        """gaussian_diffusion.py:388-429: dict(sample, pred_xstart); ``noise`` (extra, optional) replaces randn_like."""
        self._check_hooks(denoised_fn, cond_fn)
        model_kwargs = model_kwargs or {}
        model_output = model(x, self._model_timesteps(t), **model_kwargs)
        assert model_output.shape == x.shape
        if noise is None:
            noise = torch.randn_like(x)  # drawn after the model call, like the reference (:422)
        tb = self._tables(x.device)
        x0 = torch.empty_like(x, dtype=torch.float32)
        sample, _, _ = ops.cfg_step(x.float().contiguous(), model_output.float().contiguous(), noise.float().contiguous(),
                                    t.long(), tb["coef"], flags=self._flags(clip_denoised), pred_xstart=x0)
        return {"sample": sample, "pred_xstart": x0}

    def p_sample_loop(self, model, shape, noise=None, clip_denoised=True, denoised_fn=None, cond_fn=None,
                      model_kwargs=None, device=None, progress=False):
        final = None
        for sample in self.p_sample_loop_progressive(model, shape, noise=noise, clip_denoised=clip_denoised,
                                                     denoised_fn=denoised_fn, cond_fn=cond_fn, model_kwargs=model_kwargs,
                                                     device=device, progress=progress):
            final = sample
        return final["sample"]

    def _loop_start(self, model, shape, noise, device):
        if device is None:
            device = next(model.parameters()).device
        assert isinstance(shape, (tuple, list))
        img = noise if noise is not None else torch.randn(*shape, device=device)
        return img.to(device), device

    def p_sample_loop_progressive(self, model, shape, noise=None, clip_denoised=True, denoised_fn=None, cond_fn=None,
                                  model_kwargs=None, device=None, progress=False):
        img, device = self._loop_start(model, shape, noise, device)
        indices = list(range(self.num_timesteps))[::-1]
        if progress:
            from tqdm.auto import tqdm
            indices = tqdm(indices)
        for i in indices:
            t = torch.full((shape[0],), i, dtype=torch.int64, device=device)
            with torch.no_grad():
                out = self.p_sample(model, img, t, clip_denoised=clip_denoised, denoised_fn=denoised_fn, cond_fn=cond_fn,
                                    model_kwargs=model_kwargs)
                yield out
                img = out["sample"]

    def ddim_sample(self, model, x, t, clip_denoised=True, denoised_fn=None, cond_fn=None, model_kwargs=None, eta=0.0,
                    noise=None):
        """gaussian_diffusion.py:525-572."""
        self._check_hooks(denoised_fn, cond_fn)
        model_kwargs = model_kwargs or {}
        model_output = model(x, self._model_timesteps(t), **model_kwargs)
        assert model_output.shape == x.shape
        if noise is None:
            noise = torch.randn_like(x)
        tb = self._tables(x.device)
        x0 = torch.empty_like(x, dtype=torch.float32)
        sample, _, _ = ops.cfg_step(x.float().contiguous(), model_output.float().contiguous(), noise.float().contiguous(),
                                    t.long(), tb["coef"], ddim=True, eta=float(eta), flags=self._flags(clip_denoised),
                                    pred_xstart=x0)
        return {"sample": sample, "pred_xstart": x0}

    def ddim_sample_loop(self, model, shape, noise=None, clip_denoised=True, denoised_fn=None, cond_fn=None,
                         model_kwargs=None, device=None, progress=False, eta=0.0):
        final = None
        for sample in self.ddim_sample_loop_progressive(model, shape, noise=noise, clip_denoised=clip_denoised,
                                                        denoised_fn=denoised_fn, cond_fn=cond_fn, model_kwargs=model_kwargs,
                                                        device=device, progress=progress, eta=eta):
            final = sample
        return final["sample"]

    def ddim_sample_loop_progressive(self, model, shape, noise=None, clip_denoised=True, denoised_fn=None, cond_fn=None,
                                     model_kwargs=None, device=None, progress=False, eta=0.0):
        img, device = self._loop_start(model, shape, noise, device)
        indices = list(range(self.num_timesteps))[::-1]
        if progress:
            from tqdm.auto import tqdm
            indices = tqdm(indices)
        for i in indices:
            t = torch.full((shape[0],), i, dtype=torch.int64, device=device)
            with torch.no_grad():
                out = self.ddim_sample(model, img, t, clip_denoised=clip_denoised, denoised_fn=denoised_fn, cond_fn=cond_fn,
                                       model_kwargs=model_kwargs, eta=eta)
                yield out
                img = out["sample"]

    # ------------------------------------------------------------------ training loss
    def training_losses(self, model, x_start, t, model_kwargs=None, noise=None):
        """MSE branch of gaussian_diffusion.py:727-808: q_sample, model call, masked squared error per valid token
        (``terms["mse"]`` / ``terms["loss"]`` have one entry per valid token, as mean_flat over ``output[mask]`` gives)."""
        if self.loss_type.is_vb():
            raise NotImplementedError("fit_b200: the KL loss path of the reference is broken (passes train=True, :754)")
        from ..train import masked_token_mse
        model_kwargs = model_kwargs or {}
        if noise is None:
            noise = torch.randn_like(x_start)
        x_t = self.q_sample(x_start, t, noise=noise)
        model_output = model(x_t, self._model_timesteps(t), **model_kwargs)
        mask = model_kwargs["mask"]
        target = x_start if self.model_mean_type == ModelMeanType.START_X else noise
        assert model_output.shape == target.shape == x_start.shape
        mse = masked_token_mse(model_output, target, mask)
        return {"mse": mse, "loss": mse}
