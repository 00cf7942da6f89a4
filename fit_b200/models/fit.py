"""Drop-in mirror of the reference's ``models.fit`` module interface on the B200-native kernels.

Same names, constructor kwargs, parameter names / shapes (so ``load_state_dict`` of a reference
checkpoint works) and call signatures as FiT/models/fit.py:
  FiT(...)                     fit.py:362-401       FiT_models, FiT_XL_2 ... FiT_S_8   fit.py:556-617
  FiT.forward(x,t,y,pos,mask,train=True)            fit.py:474-518
  FiT.forward_with_cfg(x,t,y,pos,mask,cfg_scale)    fit.py:537-553
  FiT.patchify / unpatchify                         fit.py:444-472
The sub-modules are parameter containers whose ``forward`` also runs on the CUDA kernels; nothing here
computes with torch ops.  The model must live on a CUDA device: there is no CPU fallback.

Extra (not in the reference): ``FiT(..., precision="bf16"|"fp32")`` / ``model.set_precision()`` selects
the tcgen05 bf16 path or the fp32 verification path.
"""
from __future__ import annotations

import math
import os
from typing import Any, Dict, Literal, Optional, Union

import torch
from torch import Tensor, nn
from torch.nn.init import constant_, normal_, xavier_uniform_

from .. import ops
from .._lib import EPI_BIAS, EPI_BIAS_SILU
from ..engine import FiTEngine

__all__ = [
    "FiT", "FiT_models",
    "FiT_XL_2", "FiT_XL_4", "FiT_XL_8", "FiT_L_2", "FiT_L_4", "FiT_L_8",
    "FiT_B_2", "FiT_B_4", "FiT_B_8", "FiT_S_2", "FiT_S_4", "FiT_S_8",
]


class Linear(nn.Module):
    """nn.Linear-compatible parameter container (``weight`` [out,in], ``bias`` [out]); forward = fp32 GEMM kernel."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        self.bias = nn.Parameter(torch.empty(out_features)) if bias else None
        bound = 1.0 / math.sqrt(in_features)
        nn.init.uniform_(self.weight, -bound, bound)
        if bias:
            nn.init.uniform_(self.bias, -bound, bound)

    def forward(self, x: Tensor, epilogue: int = EPI_BIAS) -> Tensor:
        x2 = x.reshape(-1, self.in_features).float().contiguous()
        out = torch.empty(x2.shape[0], self.out_features, dtype=torch.float32, device=x.device)
        ops.gemm(x2, self.weight.detach(), None if self.bias is None else self.bias.detach(), out, epilogue)
        return out.view(*x.shape[:-1], self.out_features)


class SiLU(nn.Module):
    """Placeholder that keeps the reference's Sequential indices (``mlp.0/.2``, ``adaLN_modulation.1``);
    the activation itself is fused into the neighbouring kernels."""

    def forward(self, x):  # pragma: no cover - never called on the product path
        raise RuntimeError("fit_b200: SiLU is fused into the kernels and is not a standalone op")


class TimestepEmbedder(nn.Module):
    """fit.py:48-85: sinusoid(256) -> Linear -> SiLU -> Linear."""

    def __init__(self, hidden_size: int, frequency_embedding_size: int = 256):
        super().__init__()
        if frequency_embedding_size != 256:
            raise NotImplementedError("fit_b200: the timestep kernels are built for frequency_embedding_size = 256 (every FiT model, fit.py:55)")
        self.mlp = nn.Sequential(Linear(frequency_embedding_size, hidden_size), SiLU(), Linear(hidden_size, hidden_size))
        self.frequency_embedding_size = frequency_embedding_size

    def forward(self, t: Tensor) -> Tensor:
        f = torch.empty(t.shape[0], 256, dtype=torch.float32, device=t.device)
        freqs = torch.exp(-math.log(10000.0) * torch.arange(128, dtype=torch.float32) / 128).to(t.device)  # fit.py:73-75
        ops.timestep_sincos(t.to(torch.int64), f, freqs)
        return self.mlp[2](self.mlp[0](f, EPI_BIAS_SILU))


class LabelEmbedder(nn.Module):
    """fit.py:88-115: embedding table with a null class for classifier-free guidance + label dropout."""

    def __init__(self, num_classes: int, hidden_size: int, dropout_prob: float):
        super().__init__()
        use_cfg_embedding = dropout_prob > 0
        self.embedding_table = nn.Embedding(num_classes + use_cfg_embedding, hidden_size)
        self.num_classes = num_classes
        self.dropout_prob = dropout_prob

    def token_drop(self, labels: Tensor, force_drop_ids: Optional[Tensor] = None) -> Tensor:
        # host-side RNG draw kept as torch.rand for stream parity with the reference (fit.py:103-108)
        if force_drop_ids is None:
            drop_ids = torch.rand(labels.shape[0], device=labels.device) < self.dropout_prob
        else:
            drop_ids = force_drop_ids == 1
        return torch.where(drop_ids, self.num_classes, labels)

    def labels_for(self, labels: Tensor, train: bool, force_drop_ids: Optional[Tensor] = None) -> Tensor:
        if (train and self.dropout_prob > 0) or (force_drop_ids is not None):
            labels = self.token_drop(labels, force_drop_ids)
        return labels


def _owner(mod: nn.Module):
    """(engine, block index) of the FiT model a sub-module belongs to (set by FiT.__init__: a plain attribute, not a registered child,
    so state_dict / .to() / named_modules never see it; deepcopy and pickle follow it like the engine's own back-reference)."""
    ref = getattr(mod, "_fit_owner", None)
    fit = ref[0] if ref is not None else None
    if fit is None:
        raise RuntimeError("fit_b200: this sub-module runs on the kernels of the FiT model that owns it; build it through FiT(...)")
    return fit.engine, ref[1]


class SelfAttention(nn.Module):
    """fit.py:191-250 (parameters ``qkv``, ``proj``)."""

    def __init__(self, dim: int, num_heads: int = 8, qkv_bias: bool = False, attn_drop: float = 0.0, proj_drop: float = 0.0,
                 apply_rotate_embed: bool = False, enable_flash_attention: bool = False) -> None:
        super().__init__()
        assert dim % num_heads == 0, "dim should be divisible by num_heads"
        if enable_flash_attention:
            # the reference branch builds nn.MultiheadAttention(embed_dim=head_dim) and asserts on 4-D input
            raise AssertionError("enable_flash_attention is a broken branch of the reference (fit.py:215-218,240-243)")
        if attn_drop != 0.0 or proj_drop != 0.0:
            raise NotImplementedError("fit_b200: attention dropout is not on the reference's configured path (p=0)")
        self.num_heads = num_heads
        self.qkv = Linear(dim, dim * 3)
        self.proj = Linear(dim, dim)
        self.apply_rotate_embed = apply_rotate_embed

    def forward(self, x: Tensor, mask: Optional[Tensor] = None, freqs_cis: Optional[Tensor] = None) -> Tensor:
        """fit.py:229-250 on caller tensors (inference only: the differentiable path is the whole-network autograd node)."""
        eng, i = _owner(self)
        return eng.run_attention(i, x, mask, freqs_cis)


class SwiGLU(nn.Module):
    """fit.py:253-285 (parameters ``fc1_g``, ``fc1_x``, ``fc2``)."""

    def __init__(self, in_features: int, hidden_features: Optional[int] = None, out_features: Optional[int] = None,
                 act_layer: Any = None, norm_layer: Any = None, bias: bool = True, drop: float = 0.0) -> None:
        super().__init__()
        out_features = out_features or in_features
        hidden_features = hidden_features or in_features
        self.hidden_features = hidden_features
        self.fc1_g = Linear(in_features, hidden_features, bias=bias)
        self.fc1_x = Linear(in_features, hidden_features, bias=bias)
        self.fc2 = Linear(hidden_features, out_features, bias=bias)

    def forward(self, x: Tensor) -> Tensor:
        """fit.py:277-285 on caller tensors (inference only)."""
        eng, i = _owner(self)
        return eng.run_ffn(i, x)


class FiTBlock(nn.Module):
    """fit.py:288-337 (adaLN-Zero block).  Runs inside FiTEngine.block."""

    def __init__(self, hidden_size: int, num_heads: int, mlp_ratio: float = 4.0, ffn: Literal["swiglu", "mlp"] = "swiglu",
                 pos: Literal["rotate", "absolute"] = "rotate", **block_kwargs: Any) -> None:
        super().__init__()
        self.attn = SelfAttention(hidden_size, num_heads=num_heads, qkv_bias=True, apply_rotate_embed=(pos == "rotate"),
                                  **block_kwargs)
        if ffn == "swiglu":
            self.ffn = SwiGLU(in_features=hidden_size, hidden_features=int(hidden_size * mlp_ratio * 2 / 3))
        elif ffn == "mlp":
            # the reference's Mlp passes in_channels= to nn.Linear and raises (fit.py:130,132)
            raise TypeError("ffn='mlp' is a broken branch of the reference: Linear.__init__() got an unexpected keyword "
                            "argument 'in_channels'")
        else:
            raise ValueError(f"Unsupported ffn `{ffn}`")
        self.adaLN_modulation = nn.Sequential(SiLU(), Linear(hidden_size, 6 * hidden_size))

    def forward(self, x: Tensor, c: Tensor, mask: Optional[Tensor] = None, freqs_cis: Optional[Tensor] = None) -> Tensor:
        """fit.py:326-337 on caller tensors: x [B, N, D], c [B, D], mask [B, N] bool, freqs_cis [1|B, N, head_dim] fp32 -> [B, N, D]
        (inference only: hooks / per-block probing; training differentiates through FiT.forward's single autograd node)."""
        eng, i = _owner(self)
        return eng.run_block(i, x, c, mask, freqs_cis)


class FinalLayer(nn.Module):
    """models/dit.py:125-142 (parameters ``linear``, ``adaLN_modulation.1``)."""

    def __init__(self, hidden_size: int, patch_size: int, out_channels: int):
        super().__init__()
        self.linear = Linear(hidden_size, patch_size * patch_size * out_channels)
        self.adaLN_modulation = nn.Sequential(SiLU(), Linear(hidden_size, 2 * hidden_size))

    def forward(self, x: Tensor, c: Tensor) -> Tensor:
        """dit.py:138-142 on caller tensors (inference only)."""
        eng, _ = _owner(self)
        return eng.run_final_layer(x, c)


class FiT(nn.Module):
    """FiT: Flexible Vision Transformer for Diffusion Model -- interface of fit.py:339-553."""

    def __init__(self, patch_size: int = 2, in_channels: int = 4, hidden_size: int = 1152, depth: int = 28, num_heads: int = 16,
                 mlp_ratio: float = 4.0, class_dropout_prob: float = 0.1, num_classes: int = 1000, learn_sigma: bool = False,
                 ffn: Literal["swiglu", "mlp"] = "swiglu", pos: Literal["rotate", "absolute"] = "rotate",
                 block_kwargs: Dict[str, Any] = {}, precision: Optional[str] = None):
        super().__init__()
        self.learn_sigma = learn_sigma
        self.in_channels = in_channels
        self.out_channels = in_channels * 2 if learn_sigma else in_channels
        self.patch_size = patch_size
        self.pos = pos
        self.hidden_size = hidden_size
        self.num_heads = num_heads
        assert pos in ["absolute", "rotate"]
        assert ffn in ["swiglu", "mlp"]
        if pos != "rotate":
            raise NotImplementedError("fit_b200: pos='absolute' is unused by every caller of the reference (FiT/train.py:227)")
        if (hidden_size // num_heads) not in (64, 72):
            raise NotImplementedError("fit_b200: head_dim must be 64 or 72 (all named FiT models)")
        self.x_embedder = Linear(self.in_channels * patch_size * patch_size, hidden_size)
        self.t_embedder = TimestepEmbedder(hidden_size)
        self.y_embedder = LabelEmbedder(num_classes, hidden_size, class_dropout_prob)
        self.blocks = nn.Sequential(*[FiTBlock(hidden_size, num_heads, mlp_ratio=mlp_ratio, ffn=ffn, pos=pos, **block_kwargs)
                                      for _ in range(depth)])
        self.final_layer = FinalLayer(hidden_size, patch_size, self.out_channels)
        self.initialize_weights()
        self._engine = FiTEngine(self, precision or os.environ.get("FITB200_PRECISION", "bf16"))
        for i, blk in enumerate(self.blocks):
            for sub in (blk, blk.attn, blk.ffn):
                sub._fit_owner = (self, i)
        self.final_layer._fit_owner = (self, depth)

    # ------------------------------------------------------------------ init (fit.py:404-430)
    def initialize_weights(self):
        for mod in self.modules():
            if isinstance(mod, Linear):
                xavier_uniform_(mod.weight)
                if mod.bias is not None:
                    constant_(mod.bias, 0)
        normal_(self.y_embedder.embedding_table.weight, std=0.02)
        normal_(self.t_embedder.mlp[0].weight, std=0.02)
        normal_(self.t_embedder.mlp[2].weight, std=0.02)
        for block in self.blocks:
            constant_(block.adaLN_modulation[-1].weight, 0)
            constant_(block.adaLN_modulation[-1].bias, 0)
        constant_(self.final_layer.adaLN_modulation[-1].weight, 0)
        constant_(self.final_layer.adaLN_modulation[-1].bias, 0)
        constant_(self.final_layer.linear.weight, 0)
        constant_(self.final_layer.linear.bias, 0)

    # ------------------------------------------------------------------ precision mode
    @property
    def precision(self) -> str:
        return self._engine.precision

    def set_precision(self, precision: str) -> "FiT":
        self._engine = FiTEngine(self, precision)
        return self

    @property
    def engine(self) -> FiTEngine:
        return self._engine

    # ------------------------------------------------------------------ layout helpers (views only)
    def unpatchify(self, x: Tensor, h: int, w: int) -> Tensor:
        """(N, T, p*p*C) -> (N, C, H, W)  -- fit.py:444-456."""
        c, p = self.out_channels, self.patch_size
        nh, nw = h // p, w // p
        x = x.reshape((x.shape[0], nh, nw, p, p, c)).permute(0, 5, 1, 3, 2, 4)
        return x.reshape((x.shape[0], c, nh * p, nw * p))

    def patchify(self, x: Tensor) -> Tensor:
        """(N, C, H, W) -> (N, T, p*p*C)  -- fit.py:466-472."""
        N, C, H, W = x.shape
        p = self.patch_size
        nh, nw = H // p, W // p
        return x.reshape(N, C, nh, p, nw, p).permute(0, 2, 4, 3, 5, 1).reshape(N, nh * nw, -1)

    # ------------------------------------------------------------------ forward
    def forward(self, x: Tensor, t: Tensor, y: Tensor, pos: Tensor, mask: Tensor, train=True) -> Tensor:
        """fit.py:474-518.  train=True: tokens (N,T,p*p*C) in and out; train=False: latents (N,C,H,W) in and out
        (patchify / unpatchify are fused into the embedder / final-layer kernels)."""
        # autograd whenever gradients are wanted -- eval-mode fine-tuning / guidance included (label dropout keys off self.training alone)
        if torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
            from ..train import fit_forward_autograd
            return fit_forward_autograd(self, x, t, y, pos, mask, train)
        y = self.y_embedder.labels_for(y, self.training)
        return self._engine.forward(x.float().contiguous(), t, y, pos, mask, train=train)

    def forward_with_cfg(self, x: Tensor, t: Tensor, y: Tensor, pos: Tensor, mask: Tensor,
                         cfg_scale: Union[float, Tensor]) -> Tensor:
        """fit.py:537-553: only the first half of x is used, run with [labels | null labels], eps = u + s (c - u)
        written to both halves."""
        n = len(x)
        half = x[: n // 2].float().contiguous()
        y = self.y_embedder.labels_for(y, self.training)
        raw = self._engine.forward(half, t, y, pos, mask, train=False, rows=2 * (n // 2))
        if self.out_channels != self.in_channels:
            raise NotImplementedError("learn_sigma=True is disabled in the reference (fit.py:360,372)")
        return ops.cfg_combine(raw, float(cfg_scale))


# ---------------------------------------------------------------------- named constructors (fit.py:556-617)
def FiT_XL_2(**kwargs):
    return FiT(depth=28, hidden_size=1152, patch_size=2, num_heads=16, **kwargs)


def FiT_XL_4(**kwargs):
    return FiT(depth=28, hidden_size=1152, patch_size=4, num_heads=16, **kwargs)


def FiT_XL_8(**kwargs):
    return FiT(depth=28, hidden_size=1152, patch_size=8, num_heads=16, **kwargs)


def FiT_L_2(**kwargs):
    return FiT(depth=24, hidden_size=1024, patch_size=2, num_heads=16, **kwargs)


def FiT_L_4(**kwargs):
    return FiT(depth=24, hidden_size=1024, patch_size=4, num_heads=16, **kwargs)


def FiT_L_8(**kwargs):
    return FiT(depth=24, hidden_size=1024, patch_size=8, num_heads=16, **kwargs)


def FiT_B_2(**kwargs):
    return FiT(depth=12, hidden_size=768, patch_size=2, num_heads=12, **kwargs)


def FiT_B_4(**kwargs):
    return FiT(depth=12, hidden_size=768, patch_size=4, num_heads=12, **kwargs)


def FiT_B_8(**kwargs):
    return FiT(depth=12, hidden_size=768, patch_size=8, num_heads=12, **kwargs)


def FiT_S_2(**kwargs):
    return FiT(depth=12, hidden_size=384, patch_size=2, num_heads=6, **kwargs)


def FiT_S_4(**kwargs):
    return FiT(depth=12, hidden_size=384, patch_size=4, num_heads=6, **kwargs)


def FiT_S_8(**kwargs):
    return FiT(depth=12, hidden_size=384, patch_size=8, num_heads=6, **kwargs)


FiT_models = {
    "FiT-XL/2": FiT_XL_2, "FiT-XL/4": FiT_XL_4, "FiT-XL/8": FiT_XL_8,
    "FiT-L/2": FiT_L_2, "FiT-L/4": FiT_L_4, "FiT-L/8": FiT_L_8,
    "FiT-B/2": FiT_B_2, "FiT-B/4": FiT_B_4, "FiT-B/8": FiT_B_8,
    "FiT-S/2": FiT_S_2, "FiT-S/4": FiT_S_4, "FiT-S/8": FiT_S_8,
}
