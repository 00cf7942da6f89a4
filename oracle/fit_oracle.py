"""fp32 CPU restatement of the FiT network (TEST INFRASTRUCTURE -- see oracle/__init__.py).

Functional: every function takes the reference ``state_dict`` (same key names and
shapes as ``FiT/models/fit.py``) plus a small config dict, so that the very same
weights can be handed to the CUDA path and to this oracle.

Parity: pinned against the unmodified reference by ``oracle/make_goldens.py`` →
``tests/golden/fit_tiny_d64.npz`` / ``fit_tiny_d72.npz`` (outputs, per-block
activations, parameter gradients), checked in ``tests/test_oracle_golden.py``.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

Tensor = torch.Tensor

# name -> (depth, hidden, heads); patch size comes from the "/p" suffix. FiT/models/fit.py:556-601
_SIZES = {"S": (12, 384, 6), "B": (12, 768, 12), "L": (24, 1024, 16), "XL": (28, 1152, 16)}


def model_config(name: str, **over) -> dict:
    """Config dict for a named constructor, e.g. "FiT-XL/2" (FiT/models/fit.py:556-617)."""
    size, p = name.replace("FiT-", "").split("/")
    depth, hidden, heads = _SIZES[size]
    cfg = dict(patch_size=int(p), in_channels=4, hidden_size=hidden, depth=depth, num_heads=heads,
               mlp_ratio=4.0, class_dropout_prob=0.1, num_classes=1000, learn_sigma=False)
    cfg.update(over)
    return cfg


def swiglu_hidden(cfg: dict) -> int:
    """int(hidden * mlp_ratio * 2 / 3)  -- FiT/models/fit.py:311-312."""
    return int(cfg["hidden_size"] * cfg["mlp_ratio"] * 2 / 3)


def param_shapes(cfg: dict) -> Dict[str, tuple]:
    """Reference parameter names and shapes (FiT/models/fit.py:387-401, dit.py:129-136)."""
    D, p, C = cfg["hidden_size"], cfg["patch_size"], cfg["in_channels"]
    Hd = swiglu_hidden(cfg)
    out_c = C * 2 if cfg.get("learn_sigma", False) else C
    n_emb = cfg["num_classes"] + (1 if cfg["class_dropout_prob"] > 0 else 0)
    s = {
        "x_embedder.weight": (D, p * p * C), "x_embedder.bias": (D,),
        "t_embedder.mlp.0.weight": (D, 256), "t_embedder.mlp.0.bias": (D,),
        "t_embedder.mlp.2.weight": (D, D), "t_embedder.mlp.2.bias": (D,),
        "y_embedder.embedding_table.weight": (n_emb, D),
    }
    for i in range(cfg["depth"]):
        b = f"blocks.{i}."
        s[b + "attn.qkv.weight"] = (3 * D, D); s[b + "attn.qkv.bias"] = (3 * D,)
        s[b + "attn.proj.weight"] = (D, D); s[b + "attn.proj.bias"] = (D,)
        s[b + "ffn.fc1_g.weight"] = (Hd, D); s[b + "ffn.fc1_g.bias"] = (Hd,)
        s[b + "ffn.fc1_x.weight"] = (Hd, D); s[b + "ffn.fc1_x.bias"] = (Hd,)
        s[b + "ffn.fc2.weight"] = (D, Hd); s[b + "ffn.fc2.bias"] = (D,)
        s[b + "adaLN_modulation.1.weight"] = (6 * D, D); s[b + "adaLN_modulation.1.bias"] = (6 * D,)
    s["final_layer.linear.weight"] = (p * p * out_c, D); s["final_layer.linear.bias"] = (p * p * out_c,)
    s["final_layer.adaLN_modulation.1.weight"] = (2 * D, D)
    s["final_layer.adaLN_modulation.1.bias"] = (2 * D,)
    return s


def random_state_dict(cfg: dict, seed: int = 0, zero_init_std: float = 0.02) -> Dict[str, Tensor]:
    """Synthetic weights with the reference's init *distributions* (FiT/models/fit.py:404-430):
    xavier-uniform Linear weights, N(0,0.02) label table and t-embedder weights.  The
    parameters the reference zero-initialises (adaLN-Zero, final linear) and all biases are
    drawn N(0, zero_init_std) instead -- with them at zero the network output is identically 0
    and parity would be vacuous."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for k, shp in param_shapes(cfg).items():
        if k.endswith("bias"):
            sd[k] = torch.randn(shp, generator=g) * zero_init_std
        elif "adaLN_modulation" in k or k == "final_layer.linear.weight":
            sd[k] = torch.randn(shp, generator=g) * zero_init_std
        elif k.startswith("t_embedder") or k.startswith("y_embedder"):
            sd[k] = torch.randn(shp, generator=g) * 0.02
        else:
            bound = math.sqrt(6.0 / (shp[0] + shp[1]))
            sd[k] = (torch.rand(shp, generator=g) * 2 - 1) * bound
    return sd


# ----------------------------------------------------------------------------- pieces

def modulate(x: Tensor, shift: Tensor, scale: Tensor) -> Tensor:
    """x*(1+scale)+shift broadcast over tokens -- FiT/models/fit.py:45-46."""
    return x * (1.0 + scale[:, None, :]) + shift[:, None, :]


def layer_norm(x: Tensor) -> Tensor:
    """LayerNorm(eps=1e-6, no affine) -- FiT/models/fit.py:303,308; dit.py:131."""
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    return (x - mu) * torch.rsqrt(var + 1e-6)


def timestep_embedding(t: Tensor, dim: int = 256, max_period: float = 10000.0) -> Tensor:
    """[cos(t f) | sin(t f)], f_i = exp(-ln(max_period) i/half) -- FiT/models/fit.py:61-80."""
    half = dim // 2
    f = torch.exp(-math.log(max_period) * torch.arange(half, dtype=torch.float32) / half).to(t.device)   # host table moved to t.device (fit.py:73-75)
    a = t.to(torch.float32)[:, None] * f[None, :]
    return torch.cat([a.cos(), a.sin()], dim=1)


def t_embed(sd, t: Tensor) -> Tensor:
    """TimestepEmbedder.forward: Linear -> SiLU -> Linear -- FiT/models/fit.py:52-85."""
    h = F.linear(timestep_embedding(t), sd["t_embedder.mlp.0.weight"], sd["t_embedder.mlp.0.bias"])
    return F.linear(F.silu(h), sd["t_embedder.mlp.2.weight"], sd["t_embedder.mlp.2.bias"])


def y_embed(sd, y: Tensor) -> Tensor:
    """LabelEmbedder.forward without dropout (eval / dropout handled by caller) -- fit.py:110-115."""
    return sd["y_embedder.embedding_table.weight"][y.long()]


def rope_rotate(x: Tensor, freqs: Tensor) -> Tensor:
    """apply_rotary_emb + complex_mult for one of q/k -- FiT/models/fit.py:144-174.

    x: (B, h, N, d); freqs: (B or 1, N, d) holding interleaved (cos, sin) pairs.
    Pair j of a head: (x[2j], x[2j+1]) * (cos_j + i sin_j)."""
    B, h, N, d = x.shape
    xr = x.to(torch.float32).reshape(B, h, N, d // 2, 2)
    fr = freqs.reshape(freqs.shape[0], 1, N, d // 2, 2)
    a, b = xr[..., 0], xr[..., 1]
    c, s = fr[..., 0], fr[..., 1]
    out = torch.stack([a * c - b * s, b * c + a * s], dim=-1)
    return out.reshape(B, h, N, d).to(x.dtype)


def attention_core(q: Tensor, k: Tensor, v: Tensor, mask: Optional[Tensor]) -> Tensor:
    """softmax(q k^T d^-1/2, masked keys -> -inf) v -- FiT/models/fit.py:176-188."""
    scale = q.shape[-1] ** -0.5
    sim = torch.matmul(q, k.transpose(-2, -1)) * scale
    if mask is not None:
        sim = sim.masked_fill(~mask[:, None, None, :], float("-inf"))
    return torch.matmul(torch.softmax(sim, dim=-1), v)


def self_attention(sd, pre: str, x: Tensor, mask, freqs, num_heads: int) -> Tensor:
    """SelfAttention.forward -- FiT/models/fit.py:229-250 (qkv column = which*D + head*d + i)."""
    B, N, D = x.shape
    qkv = F.linear(x, sd[pre + "attn.qkv.weight"], sd[pre + "attn.qkv.bias"])
    qkv = qkv.reshape(B, N, 3, num_heads, D // num_heads).permute(2, 0, 3, 1, 4)
    q, k, v = qkv[0], qkv[1], qkv[2]
    if freqs is not None:
        q, k = rope_rotate(q, freqs), rope_rotate(k, freqs)
    o = attention_core(q, k, v, mask)
    o = o.transpose(1, 2).reshape(B, N, D)
    return F.linear(o, sd[pre + "attn.proj.weight"], sd[pre + "attn.proj.bias"])


def swiglu(sd, pre: str, x: Tensor) -> Tensor:
    """fc2(silu(fc1_g x) * fc1_x x) -- FiT/models/fit.py:277-285."""
    g = F.linear(x, sd[pre + "ffn.fc1_g.weight"], sd[pre + "ffn.fc1_g.bias"])
    u = F.linear(x, sd[pre + "ffn.fc1_x.weight"], sd[pre + "ffn.fc1_x.bias"])
    return F.linear(F.silu(g) * u, sd[pre + "ffn.fc2.weight"], sd[pre + "ffn.fc2.bias"])


def block_forward(sd, i: int, x: Tensor, c: Tensor, mask, freqs, num_heads: int) -> Tensor:
    """FiTBlock.forward (adaLN-Zero) -- FiT/models/fit.py:326-337."""
    pre = f"blocks.{i}."
    mod = F.linear(F.silu(c), sd[pre + "adaLN_modulation.1.weight"], sd[pre + "adaLN_modulation.1.bias"])
    sh1, sc1, g1, sh2, sc2, g2 = mod.chunk(6, dim=1)
    x = x + g1[:, None, :] * self_attention(sd, pre, modulate(layer_norm(x), sh1, sc1), mask, freqs, num_heads)
    x = x + g2[:, None, :] * swiglu(sd, pre, modulate(layer_norm(x), sh2, sc2))
    return x


def final_layer(sd, x: Tensor, c: Tensor) -> Tensor:
    """FinalLayer.forward -- FiT/models/dit.py:138-142."""
    mod = F.linear(F.silu(c), sd["final_layer.adaLN_modulation.1.weight"],
                   sd["final_layer.adaLN_modulation.1.bias"])
    shift, scale = mod.chunk(2, dim=1)
    return F.linear(modulate(layer_norm(x), shift, scale), sd["final_layer.linear.weight"],
                    sd["final_layer.linear.bias"])


def patchify(x: Tensor, p: int) -> Tensor:
    """(N,C,H,W) -> (N, nh*nw, p*p*C), inner order (ph, pw, c) -- FiT/models/fit.py:466-472."""
    N, C, H, W = x.shape
    nh, nw = H // p, W // p
    return x.reshape(N, C, nh, p, nw, p).permute(0, 2, 4, 3, 5, 1).reshape(N, nh * nw, p * p * C)


def unpatchify(x: Tensor, h: int, w: int, p: int, c: int) -> Tensor:
    """(N,T,p*p*C) -> (N,C,H,W) -- FiT/models/fit.py:444-456."""
    nh, nw = h // p, w // p
    x = x.reshape(x.shape[0], nh, nw, p, p, c).permute(0, 5, 1, 3, 2, 4)
    return x.reshape(x.shape[0], c, nh * p, nw * p)


# ----------------------------------------------------------------------------- network

def forward(sd, cfg: dict, x: Tensor, t: Tensor, y: Tensor, pos: Tensor, mask: Tensor,
            train: bool = True, collect: Optional[List[Tensor]] = None) -> Tensor:
    """FiT.forward with pos="rotate" and eval-mode label embedder -- FiT/models/fit.py:474-518.

    ``collect`` (optional list) receives the token stream after the embedder and after every
    block, for per-block parity checks."""
    p, D = cfg["patch_size"], cfg["hidden_size"]
    if not train:
        _, _, h, w = x.shape
        x = patchify(x, p)
    x = F.linear(x, sd["x_embedder.weight"], sd["x_embedder.bias"])
    c = t_embed(sd, t) + y_embed(sd, y)
    if collect is not None:
        collect.append(x)
    for i in range(cfg["depth"]):
        x = block_forward(sd, i, x, c, mask, pos, cfg["num_heads"])
        if collect is not None:
            collect.append(x)
    x = final_layer(sd, x, c)
    if not train:
        x = unpatchify(x, h, w, p, cfg["in_channels"])
    return x


def forward_with_cfg(sd, cfg: dict, x, t, y, pos, mask, cfg_scale: float) -> Tensor:
    """FiT.forward_with_cfg -- FiT/models/fit.py:537-553 (learn_sigma=False: `rest` is empty)."""
    half = x[: len(x) // 2]
    out = forward(sd, cfg, torch.cat([half, half], 0), t, y, pos, mask, train=False)
    C = cfg["in_channels"]
    eps, rest = out[:, :C], out[:, C:]
    cond, uncond = torch.split(eps, len(eps) // 2, dim=0)
    g = uncond + cfg_scale * (cond - uncond)
    return torch.cat([torch.cat([g, g], 0), rest], dim=1)


def masked_mse(out: Tensor, target: Tensor, mask: Tensor) -> Tensor:
    """F.mse_loss(out[mask], target[mask]) -- FiT/train.py:63-66, masked_FiT/train.py:55-58."""
    return F.mse_loss(out[mask], target[mask])


def loss_and_grads(sd, cfg, x_t, t, y, pos, mask, target):
    """Training-step arithmetic through autograd of this restatement (FiT/train.py:55-66)."""
    leaf = {k: v.detach().clone().requires_grad_(True) for k, v in sd.items()}
    out = forward(leaf, cfg, x_t, t, y, pos, mask, train=True)
    loss = masked_mse(out, target, mask)
    loss.backward()
    return loss.detach(), out.detach(), {k: v.grad for k, v in leaf.items()}


# ----------------------------------------------------------------------------- sampling helpers
# (FiT/train.py:97-166: the five padding helpers the sampling script calls)

def pad_latent(x: Tensor, p: int, max_size: int, max_length: int) -> Tensor:
    """_pad_latent: tokens zero-padded to max_length, re-laid as a max_size^2 latent
    (FiT/train.py:115-126); returned unchanged if it has more than max_length tokens."""
    n, c, _, _ = x.shape
    tok = patchify(x, p)
    if tok.shape[1] > max_length:
        return x
    buf = torch.zeros((n, max_length, p * p * c), dtype=x.dtype)
    buf[:, : tok.shape[1]] = tok
    return unpatchify(buf, max_size, max_size, p, c)


def unpad_latent(x: Tensor, valid_t: int, h: int, w: int, p: int) -> Tensor:
    """_unpad_latent -- FiT/train.py:128-134."""
    c = x.shape[1]
    return unpatchify(patchify(x, p)[:, :valid_t], h, w, p, c)


def create_pos_embed(h: int, w: int, p: int, max_length: int, dim: int):
    """_create_pos_embed (rotate, VisionNTK always on) -- FiT/train.py:136-156."""
    from .pos_embed_oracle import precompute_freqs_cis_2d
    fill = torch.tensor(precompute_freqs_cis_2d(dim, h // p, w // p, max_length=max_length), dtype=torch.float32)
    if fill.shape[0] > max_length:
        pe = fill
    else:
        pe = torch.zeros((max_length, dim), dtype=torch.float32)
        pe[: fill.shape[0]] = fill
    return pe[None], fill.shape[0]


def create_mask(valid_t: int, max_length: int, n: int) -> Tensor:
    """_create_mask -- FiT/train.py:158-166."""
    if valid_t > max_length:
        m = torch.ones((valid_t,), dtype=torch.bool)
    else:
        m = torch.zeros((max_length,), dtype=torch.bool)
        m[:valid_t] = True
    return m[None].repeat(n, 1)
