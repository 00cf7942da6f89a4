"""Forward executor of the FiT network on libfit_b200.so.

Sequences the C-ABI kernels for FiT.forward (reference FiT/models/fit.py:474-518): no torch
arithmetic, no allocation after the first call for a given (rows, tokens) shape, everything on the
current stream -- so a whole forward (and a whole sampling step) is CUDA-graph capturable.

Data layout in HBM (row-major, one row per token):
  xs   [R*T, D]   residual stream (act dtype)      h   [R*T, D]   LN+modulate output (GEMM A operand)
  qkv  [R*T, 3D]  column = which*D + head*d + i    att [R*T, D]   attention output, heads merged
  g    [R*T, Hp]  SwiGLU hidden                     mod [R, L*6D+2D] fp32 adaLN table of ALL blocks + final layer
Precision modes: "bf16" (tcgen05 GEMMs, bf16 activations, fp32 accumulation / LN / softmax / modulation)
and "fp32" (verification mode: CUDA-core fp32 GEMMs and fp32 activations).
"""
from __future__ import annotations

import math
import os
from typing import Dict, Optional, Tuple

import torch

from . import ops
from ._lib import EPI_BIAS, EPI_BIAS_GATE_RESID, EPI_BIAS_ROPE, EPI_BIAS_SILU, EPI_SWIGLU
from ._nvtx import rng

Tensor = torch.Tensor


def _round_up(a: int, b: int) -> int:
    return (a + b - 1) // b * b


class _Weights:
    """Device copies of the parameters in the layout the kernels want (rebuilt when parameters change)."""

    def __init__(self, model, precision: str):
        D, L = model.hidden_size, len(model.blocks)
        H = model.blocks[0].ffn.hidden_features if L else 0
        Hp = _round_up(H, 16)
        self.H, self.Hp = H, Hp
        bf16 = precision == "bf16"
        wdt = torch.bfloat16 if bf16 else torch.float32

        def f32(p):
            return p.detach().float().contiguous()

        def w(p):
            return p.detach().to(wdt).contiguous()

        self.x_w, self.x_b = f32(model.x_embedder.weight), f32(model.x_embedder.bias)
        self.t0_w, self.t0_b = f32(model.t_embedder.mlp[0].weight), f32(model.t_embedder.mlp[0].bias)
        self.t2_w, self.t2_b = f32(model.t_embedder.mlp[2].weight), f32(model.t_embedder.mlp[2].bias)
        self.y_table = f32(model.y_embedder.embedding_table.weight)
        # fit.py:73-75: the sinusoid frequencies are a host-side fp32 table in the reference; computing them the
        # same way (CPU torch.exp) keeps the arguments t*f bit-identical to the reference's
        half = 128
        self.t_freqs = torch.exp(-math.log(10000.0) * torch.arange(half, dtype=torch.float32) / half).to(self.y_table.device)
        # adaLN of every block and of the final layer as ONE [L*6D + 2D, D] weight: c is shared by all of
        # them (fit.py:505,513; dit.py:139), so a single small-M GEMM per forward produces the whole table
        ada_w = [b.adaLN_modulation[1].weight for b in model.blocks] + [model.final_layer.adaLN_modulation[1].weight]
        ada_b = [b.adaLN_modulation[1].bias for b in model.blocks] + [model.final_layer.adaLN_modulation[1].bias]
        self.ada_w = torch.cat([p.detach() for p in ada_w], 0).to(wdt).contiguous()
        self.ada_b = torch.cat([p.detach() for p in ada_b], 0).float().contiguous()
        self.n_mod = self.ada_w.shape[0]
        self.blocks = []
        for blk in model.blocks:
            e = {}
            e["qkv_w"], e["qkv_b"] = w(blk.attn.qkv.weight), f32(blk.attn.qkv.bias)
            e["proj_w"], e["proj_b"] = w(blk.attn.proj.weight), f32(blk.attn.proj.bias)
            fg, fx = blk.ffn.fc1_g, blk.ffn.fc1_x
            if bf16:
                # packed SwiGLU weight: 16 rows of fc1_g then the matching 16 rows of fc1_x, so one 32-column
                # accumulator chunk of the epilogue holds gate and up of the same 16 hidden units
                dev = fg.weight.device
                wg = torch.zeros(Hp, D, device=dev, dtype=wdt)
                wx = torch.zeros(Hp, D, device=dev, dtype=wdt)
                wg[:H], wx[:H] = fg.weight.detach().to(wdt), fx.weight.detach().to(wdt)
                bg = torch.zeros(Hp, device=dev)
                bx = torch.zeros(Hp, device=dev)
                bg[:H], bx[:H] = fg.bias.detach().float(), fx.bias.detach().float()
                e["fc1_w"] = torch.stack([wg.view(Hp // 16, 16, D), wx.view(Hp // 16, 16, D)], 1).reshape(2 * Hp, D).contiguous()
                e["fc1_b"] = torch.stack([bg.view(Hp // 16, 16), bx.view(Hp // 16, 16)], 1).reshape(2 * Hp).contiguous()
            else:
                e["fc1_g_w"], e["fc1_g_b"] = f32(fg.weight), f32(fg.bias)
                e["fc1_x_w"], e["fc1_x_b"] = f32(fx.weight), f32(fx.bias)
            w2 = blk.ffn.fc2.weight.detach()
            if Hp != H:
                w2p = torch.zeros(D, Hp, device=w2.device, dtype=wdt)
                w2p[:, :H] = w2.to(wdt)
                e["fc2_w"] = w2p
            else:
                e["fc2_w"] = w2.to(wdt).contiguous()
            e["fc2_b"] = f32(blk.ffn.fc2.bias)
            self.blocks.append(e)
        self.fin_w, self.fin_b = f32(model.final_layer.linear.weight), f32(model.final_layer.linear.bias)
        self.fin_w16 = w(model.final_layer.linear.weight) if bf16 else None   # tensor-core final layer (GEMM with the LayerNorm prologue)


LN_EPS = 1e-6  # norm1 / norm2 (fit.py:303,308)


def ln_prologue_ok(precision: str, M: int, D: int, T: int) -> bool:
    """Whether FiTEngine runs modulate(norm(x)) as the PROLOGUE of the qkv / fc1 GEMMs (gemm_tc.cu, LNA) instead of the stand-alone
    ln_modulate kernel.  Opt-in (FITB200_LN_FUSE=1): parity-green, but measured SLOWER on the headline step (128.6 vs 122.6 ms,
    profiles/r2b_ln_prologue.txt) -- the 256 x 256 pair-tile mainloop already sits at the shared-memory / L2->SM bandwidth edge, and rescaling
    the operand on its way to the tensor core costs it more (65-70 % tensor pipe instead of 80 %) than the separate 0.13 ms pass over HBM.
    Shapes served: bf16, every 128-row CTA tile inside one sample (T % 128 == 0: the reference's 256-token sequences and the 1024 / 4096-token
    extrapolation shapes), width a whole number of 64-column k-slabs (every FiT size: 384, 768, 1024, 1152)."""
    return precision == "bf16" and M > 128 and T % 128 == 0 and D > 128 and D % 64 == 0 and os.environ.get("FITB200_LN_FUSE", "0") == "1"


def final_gemm_ok(precision: str, M: int, D: int, T: int) -> bool:
    """The final layer (LayerNorm + modulate + Linear(D -> p*p*C), dit.py:138-142) as ONE tcgen05 GEMM with the LayerNorm prologue: with 16
    output columns the GEMM is a single pass over the residual stream, so here the prologue is a clear gain (0.1 ms against 0.6 ms for the
    CUDA-core kernel, which issues 576 fp32 FMAs per token and lane).  Same shape conditions as the prologue; FITB200_FINAL_GEMM=0 keeps the
    CUDA-core kernel (the fp32 verification mode and every other shape take it anyway)."""
    return precision == "bf16" and M > 128 and T % 128 == 0 and D > 128 and D % 64 == 0 and os.environ.get("FITB200_FINAL_GEMM", "1") != "0"


def ln_stats_ok(precision: str, D: int) -> bool:
    """bf16 path: the proj / fc2 GEMM epilogues emit the row sums of the residual stream they write (stats_out), so modulate(norm(x)) is a
    streaming pass (ln_apply) instead of a row-reduction kernel.  FITB200_LN_STATS=0 keeps ln_modulate (A/B)."""
    return precision == "bf16" and D % 8 == 0 and os.environ.get("FITB200_LN_STATS", "1") != "0"


class _Workspace:
    def __init__(self, R: int, T: int, D: int, Hp: int, n_mod: int, act: torch.dtype, device):
        M = R * T
        z = dict(device=device)
        self.R, self.T = R, T
        L6 = (n_mod - 2 * D) // (6 * D)
        self.modA = None
        if ln_prologue_ok("bf16" if act == torch.bfloat16 else "fp32", M, D, T) and L6 > 0:
            # LN prologue operands: packed modulation rows of every block's norm1 / norm2, the (c1, c2) vectors of the current qkv / fc1 GEMM,
            # partial row sums of the residual stream
            self.modA = torch.empty(2 * L6, 2 * R, D, dtype=torch.bfloat16, **z)
            self.cq = torch.empty(2 * R, 3 * D, dtype=torch.float32, **z)
            self.cf = torch.empty(2 * R, 2 * Hp, dtype=torch.float32, **z)
        self.parts = None
        if self.modA is not None or ln_stats_ok("bf16" if act == torch.bfloat16 else "fp32", D):
            self.parts = torch.empty(M, ops.gemm_stats_parts(D), 2, dtype=torch.float32, **z)   # partial row sums of the residual stream
        # final layer as a GEMM with the LayerNorm + modulate PROLOGUE (dit.py:138-142): packed modulation rows, (c1, c2) vectors, token output
        self.fin_mod = None
        if final_gemm_ok("bf16" if act == torch.bfloat16 else "fp32", M, D, T) and self.parts is not None:
            self.fin_mod = torch.empty(1, 2 * R, D, dtype=torch.bfloat16, **z)
            self.fin_c = None      # [2R, P] fp32, allocated at first use (P is the caller's)
            self.fin_tok = None    # [M, P] fp32
        self.xs = torch.empty(M, D, dtype=act, **z)
        self.h = torch.empty(M, D, dtype=act, **z)
        self.qkv = torch.empty(M, 3 * D, dtype=act, **z)
        self.att = torch.empty(M, D, dtype=act, **z)
        self.g = torch.zeros(M, Hp, dtype=act, **z)  # padded columns (FiT-L) stay zero
        self.mod = torch.empty(R, n_mod, dtype=torch.float32, **z)
        self.tfreq = torch.empty(R, 256, dtype=torch.float32, **z)
        self.t1 = torch.empty(R, D, dtype=torch.float32, **z)
        self.t2 = torch.empty(R, D, dtype=torch.float32, **z)
        self.c = torch.empty(R, D, dtype=torch.float32, **z)
        self.silu_c = torch.empty(R, D, dtype=act, **z)
        self.kv_len = torch.empty(R, dtype=torch.int32, **z)
        self.n_valid = torch.empty(R, dtype=torch.int32, **z)
        # per-block views into the adaLN table: shift_msa, scale_msa, gate_msa, shift_mlp, scale_mlp, gate_mlp
        # (chunk order of fit.py:329-330); final layer: shift, scale (dit.py:139)
        self.mod_views = [[self.mod[:, (6 * i + k) * D:(6 * i + k + 1) * D] for k in range(6)] for i in range(L6)]
        self.fin_views = [self.mod[:, 6 * L6 * D + k * D: 6 * L6 * D + (k + 1) * D] for k in range(2)]


class FiTEngine:
    """Runs FiT.forward for one ``fit_b200.models.fit.FiT`` module."""

    def __init__(self, model, precision: str = "bf16"):
        if precision not in ("bf16", "fp32"):
            raise ValueError("precision must be 'bf16' or 'fp32'")
        self.model = model
        self.precision = precision
        self._w: Optional[_Weights] = None
        self._w_key = None
        self._ws: Dict[Tuple[int, int, str], _Workspace] = {}

    # ------------------------------------------------------------------ caches
    @property
    def act_dtype(self):
        return torch.bfloat16 if self.precision == "bf16" else torch.float32

    def weights(self) -> _Weights:
        params = list(self.model.parameters())
        key = (self.precision,) + tuple((p.data_ptr(), p._version) for p in params)
        if self._w is None or key != self._w_key:
            if not params[0].is_cuda:
                raise RuntimeError("fit_b200: the model must live on a CUDA device (no CPU fallback)")
            self._w = _Weights(self.model, self.precision)
            self._w_key = key
        return self._w

    def workspace(self, R: int, T: int, device) -> _Workspace:
        key = (R, T, str(device))
        ws = self._ws.get(key)
        if ws is None:
            W = self.weights()
            if len(self._ws) >= 4:  # bounded cache: mixed-resolution callers cycle through few shapes
                self._ws.pop(next(iter(self._ws)))
            ws = _Workspace(R, T, self.model.hidden_size, W.Hp, W.n_mod, self.act_dtype, device)
            self._ws[key] = ws
        return ws

    # ------------------------------------------------------------------ pieces
    def conditioning(self, ws: _Workspace, W: _Weights, t: Tensor, y: Tensor):
        """t_embedder + y_embedder + c = t + y (fit.py:503-505) and the adaLN table of all blocks."""
        ops.timestep_sincos(t, ws.tfreq, W.t_freqs)
        ops.gemm(ws.tfreq, W.t0_w, W.t0_b, ws.t1, EPI_BIAS_SILU)
        ops.gemm(ws.t1, W.t2_w, W.t2_b, ws.t2, EPI_BIAS)
        ops.cond_combine(ws.t2, y, W.y_table, ws.c, ws.silu_c)
        ops.gemm(ws.silu_c, W.ada_w, W.ada_b, ws.mod, EPI_BIAS)
        if getattr(ws, "modA", None) is not None:
            D = self.model.hidden_size
            ops.modulation_pack(ws.mod, ws.mod[:, D:], ws.mod.stride(0), ws.R, D, ws.modA.shape[0], 3 * D, ws.modA)

    def block(self, i: int, ws: _Workspace, W: _Weights, pos: Tensor, mask: Optional[Tensor]):
        """FiTBlock.forward (fit.py:326-337) in place on ws.xs."""
        m = self.model
        R, T, D = ws.R, ws.T, m.hidden_size
        e = W.blocks[i]
        sh1, sc1, g1, sh2, sc2, g2 = ws.mod_views[i]
        nm = ws.mod.stride(0)
        if getattr(ws, "modA", None) is not None:
            # modulate(norm1(x)) and modulate(norm2(x)) (fit.py:332,336) as the PROLOGUE of the qkv / fc1 GEMMs: the modulated rows never exist
            # in HBM; the row sums of the residual stream come out of the proj / fc2 epilogues (ws.parts)
            ops.gemm(ws.modA[2 * i], e["qkv_w"], e["qkv_b"], ws.cq, EPI_BIAS, bias_row0=R)
            ops.gemm(ws.xs, e["qkv_w"], None, ws.qkv, EPI_BIAS_ROPE, rope=pos, rope_cols=2 * D, rows_per_sample=T,
                     ln=(ws.modA[2 * i, :R], ws.cq, ws.parts, LN_EPS))
            ops.attention(ws.qkv, None, ws.kv_len, mask, R, T, m.num_heads, D // m.num_heads, ws.att, prerotated=True)
            ops.gemm(ws.att, e["proj_w"], e["proj_b"], ws.xs, EPI_BIAS_GATE_RESID, gate=g1, ld_gate=nm, rows_per_sample=T, resid=ws.xs,
                     stats_out=ws.parts)
            ops.gemm(ws.modA[2 * i + 1], e["fc1_w"], e["fc1_b"], ws.cf, EPI_BIAS, bias_row0=R)
            ops.gemm(ws.xs, e["fc1_w"], None, ws.g, EPI_SWIGLU, rows_per_sample=T, ln=(ws.modA[2 * i + 1, :R], ws.cf, ws.parts, LN_EPS))
            ops.gemm(ws.g, e["fc2_w"], e["fc2_b"], ws.xs, EPI_BIAS_GATE_RESID, gate=g2, ld_gate=nm, rows_per_sample=T, resid=ws.xs,
                     stats_out=ws.parts)
            return
        parts = getattr(ws, "parts", None)   # row sums of ws.xs (from row_stats / the previous GATE_RESID epilogue): LN without the row reduction
        if parts is not None:
            ops.ln_apply(ws.xs, parts, sh1, sc1, nm, T, ws.h, LN_EPS)
        else:
            ops.ln_modulate(ws.xs, sh1, sc1, nm, T, ws.h)
        if self.precision == "bf16":
            # apply_rotary_emb (fit.py:160-174) rides in the qkv GEMM epilogue (fp32 accumulators, q and k columns), so the
            # attention kernel fetches finished operands with TMA
            ops.gemm(ws.h, e["qkv_w"], e["qkv_b"], ws.qkv, EPI_BIAS_ROPE, rope=pos, rope_cols=2 * D)
            ops.attention(ws.qkv, None, ws.kv_len, mask, R, T, m.num_heads, D // m.num_heads, ws.att, prerotated=True)
        else:
            ops.gemm(ws.h, e["qkv_w"], e["qkv_b"], ws.qkv, EPI_BIAS)
            ops.attention(ws.qkv, pos, ws.kv_len, mask, R, T, m.num_heads, D // m.num_heads, ws.att)
        ops.gemm(ws.att, e["proj_w"], e["proj_b"], ws.xs, EPI_BIAS_GATE_RESID, gate=g1, ld_gate=nm, rows_per_sample=T, resid=ws.xs, stats_out=parts)
        if parts is not None:
            ops.ln_apply(ws.xs, parts, sh2, sc2, nm, T, ws.h, LN_EPS)
        else:
            ops.ln_modulate(ws.xs, sh2, sc2, nm, T, ws.h)
        if self.precision == "bf16":
            ops.gemm(ws.h, e["fc1_w"], e["fc1_b"], ws.g, EPI_SWIGLU)
        else:
            ops.gemm(ws.h, e["fc1_g_w"], e["fc1_g_b"], ws.g, EPI_SWIGLU, B2=e["fc1_x_w"], bias2=e["fc1_x_b"])
        ops.gemm(ws.g, e["fc2_w"], e["fc2_b"], ws.xs, EPI_BIAS_GATE_RESID, gate=g2, ld_gate=nm, rows_per_sample=T, resid=ws.xs, stats_out=parts)

    # ------------------------------------------------------------------ sub-module entry points (fit.py:229-250,277-285,326-337; dit.py:138-142)
    class _Scratch:
        """Buffers of one stand-alone sub-module call (same fields FiTEngine.block reads from a _Workspace)."""
        pass

    def _scratch(self, x: Tensor, mask: Optional[Tensor], Hp: int):
        R, T, D = x.shape
        M, act, dev = R * T, self.act_dtype, x.device
        ws = FiTEngine._Scratch()
        ws.R, ws.T = R, T
        ws.xs = ops.cast(x.reshape(M, D).contiguous(), torch.empty(M, D, dtype=act, device=dev)) if x.dtype != act else x.reshape(M, D).clone()
        ws.h = torch.empty(M, D, dtype=act, device=dev)
        ws.qkv = torch.empty(M, 3 * D, dtype=act, device=dev)
        ws.att = torch.empty(M, D, dtype=act, device=dev)
        ws.g = torch.zeros(M, Hp, dtype=act, device=dev)
        if mask is None:
            mask = torch.ones(R, T, dtype=torch.bool, device=dev)
        ws.mask = mask
        ws.kv_len, ws.n_valid = ops.mask_to_kvlen(mask)
        return ws

    def _check_sub(self, x: Tensor):
        if x.dim() != 3 or x.shape[2] != self.model.hidden_size:
            raise ValueError(f"expected x of shape (batch, tokens, {self.model.hidden_size}), got {tuple(x.shape)}")
        if x.dtype not in (torch.float32, torch.bfloat16):
            raise TypeError("fit_b200: x must be fp32 or bf16")
        if not x.is_cuda:
            raise RuntimeError("fit_b200 kernels need CUDA tensors (there is no CPU fallback)")

    def _modulation(self, c: Tensor, w: Tensor, b: Tensor) -> Tensor:
        """adaLN_modulation(c) = Linear(SiLU(c)) (fit.py:323-324,329; dit.py:134-139) for ONE block / the final layer -> fp32 [R, n]."""
        c32 = c.float().contiguous()
        sc = ops.silu_fwd(c32, torch.empty_like(c32))
        if self.act_dtype != torch.float32:
            sc = ops.cast(sc, torch.empty(c32.shape, dtype=self.act_dtype, device=c.device))
        return ops.gemm(sc, w, b, torch.empty(c.shape[0], w.shape[0], dtype=torch.float32, device=c.device), EPI_BIAS)

    def _pos(self, freqs_cis: Optional[Tensor], R: int, T: int) -> Tensor:
        d = self.model.hidden_size // self.model.num_heads
        if freqs_cis is None:
            raise ValueError("fit_b200: freqs_cis (the RoPE table) is required: every FiT block is built with pos='rotate'")
        pos = freqs_cis if freqs_cis.dim() == 3 else freqs_cis[None]
        if pos.shape[0] not in (1, R) or pos.shape[1] != T or pos.shape[2] != d or pos.dtype != torch.float32:
            raise ValueError(f"freqs_cis {tuple(pos.shape)} {pos.dtype} does not match (1|{R}, {T}, {d}) fp32")
        return pos.contiguous()

    def _out(self, buf: Tensor, like: Tensor, shape) -> Tensor:
        if buf.dtype != like.dtype:
            buf = ops.cast(buf, torch.empty(buf.shape, dtype=like.dtype, device=buf.device))
        return buf.view(*shape)

    def run_block(self, i: int, x: Tensor, c: Tensor, mask: Optional[Tensor], freqs_cis: Optional[Tensor]) -> Tensor:
        """FiTBlock.forward(x, c, mask, freqs_cis) (fit.py:326-337) on caller tensors: same kernels as the whole-network path."""
        self._check_sub(x)
        W = self.weights()
        D = self.model.hidden_size
        R, T, _ = x.shape
        ws = self._scratch(x, mask, W.Hp)
        ws.mod = self._modulation(c, W.ada_w[6 * i * D:6 * (i + 1) * D], W.ada_b[6 * i * D:6 * (i + 1) * D])
        ws.mod_views = {i: [ws.mod[:, k * D:(k + 1) * D] for k in range(6)]}
        self.block(i, ws, W, self._pos(freqs_cis, R, T), ws.mask)
        return self._out(ws.xs, x, x.shape)

    def run_attention(self, i: int, x: Tensor, mask: Optional[Tensor], freqs_cis: Optional[Tensor]) -> Tensor:
        """SelfAttention.forward(x, mask, freqs_cis) (fit.py:229-250): qkv GEMM (+RoPE), attention, proj GEMM (+bias)."""
        self._check_sub(x)
        W, m = self.weights(), self.model
        e = W.blocks[i]
        R, T, D = x.shape
        ws = self._scratch(x, mask, 16)
        pos = self._pos(freqs_cis, R, T)
        h, d = m.num_heads, D // m.num_heads
        if self.precision == "bf16":
            ops.gemm(ws.xs, e["qkv_w"], e["qkv_b"], ws.qkv, EPI_BIAS_ROPE, rope=pos, rope_cols=2 * D)
            ops.attention(ws.qkv, None, ws.kv_len, ws.mask, R, T, h, d, ws.att, prerotated=True)
        else:
            ops.gemm(ws.xs, e["qkv_w"], e["qkv_b"], ws.qkv, EPI_BIAS)
            ops.attention(ws.qkv, pos, ws.kv_len, ws.mask, R, T, h, d, ws.att)
        ops.gemm(ws.att, e["proj_w"], e["proj_b"], ws.h, EPI_BIAS)
        return self._out(ws.h, x, x.shape)

    def run_ffn(self, i: int, x: Tensor) -> Tensor:
        """SwiGLU.forward(x) (fit.py:277-285): fc2(silu(fc1_g(x)) * fc1_x(x))."""
        self._check_sub(x)
        W = self.weights()
        e = W.blocks[i]
        ws = self._scratch(x, None, W.Hp)
        if self.precision == "bf16":
            ops.gemm(ws.xs, e["fc1_w"], e["fc1_b"], ws.g, EPI_SWIGLU)
        else:
            ops.gemm(ws.xs, e["fc1_g_w"], e["fc1_g_b"], ws.g, EPI_SWIGLU, B2=e["fc1_x_w"], bias2=e["fc1_x_b"])
        ops.gemm(ws.g, e["fc2_w"], e["fc2_b"], ws.h, EPI_BIAS)
        return self._out(ws.h, x, x.shape)

    def run_final_layer(self, x: Tensor, c: Tensor) -> Tensor:
        """FinalLayer.forward(x, c) (dit.py:138-142): tokens [R, T, D] -> [R, T, p*p*C_out]."""
        self._check_sub(x)
        W, m = self.weights(), self.model
        R, T, D = x.shape
        L = len(m.blocks)
        ws = self._scratch(x, None, 16)
        mod = self._modulation(c, W.ada_w[6 * L * D:], W.ada_b[6 * L * D:])
        P = m.patch_size * m.patch_size * m.out_channels
        out = torch.empty(R, T, P, dtype=torch.float32, device=x.device)
        ops.final_layer(ws.xs, mod[:, :D], mod[:, D:], mod.stride(0), W.fin_w, W.fin_b, out, R, T, m.patch_size, m.out_channels)
        return out if x.dtype == torch.float32 else out.to(x.dtype)

    # ------------------------------------------------------------------ forward
    def forward(self, x: Tensor, t: Tensor, y: Tensor, pos: Tensor, mask: Tensor, train: bool = True,
                rows: Optional[int] = None, out: Optional[Tensor] = None, collect: Optional[list] = None) -> Tensor:
        """x: tokens [N,T,p*p*C] (train=True) or latents [N,C,H,W] (train=False), fp32.
        rows: number of rows to run (>= x.shape[0], a multiple of it): row r reads x[r % x.shape[0]] --
        forward_with_cfg's cat([half, half]) without the copy."""
        m = self.model
        p, C_in, D = m.patch_size, m.in_channels, m.hidden_size
        if x.dtype != torch.float32:
            raise TypeError("fit_b200: x must be fp32")
        R = x.shape[0] if rows is None else rows
        if train:
            T = x.shape[1]
            H = Wd = 0
        else:
            H, Wd = x.shape[2], x.shape[3]
            T = (H // p) * (Wd // p)
        if mask.shape[0] != R or mask.shape[1] != T:
            raise ValueError(f"mask shape {tuple(mask.shape)} does not match ({R}, {T})")
        if pos.dim() == 2:
            pos = pos[None]
        if pos.shape[0] not in (1, R) or pos.shape[1] != T or pos.shape[2] != D // m.num_heads:
            raise ValueError(f"pos shape {tuple(pos.shape)} does not match (1|{R}, {T}, {D // m.num_heads})")
        if pos.dtype != torch.float32:
            raise TypeError("fit_b200: pos must be fp32 (cos, sin) pairs")
        pos = pos.contiguous()
        W = self.weights()
        ws = self.workspace(R, T, x.device)
        t = t.to(torch.int64) if t.dtype != torch.int64 else t
        y = y.to(torch.int64) if y.dtype != torch.int64 else y
        with rng("fit/embed+conditioning"):
            ops.mask_to_kvlen(mask, ws.kv_len, ws.n_valid)
            ops.embed_tokens(x, W.x_w, W.x_b, ws.xs, R, T, p, C_in)
            self.conditioning(ws, W, t, y)
            if ws.parts is not None:
                ops.row_stats(ws.xs, ws.parts)
        if collect is not None:
            collect.append(ws.xs.float().view(R, T, D).clone())
        for i in range(len(m.blocks)):
            with rng(f"fit/block/{i}"):
                self.block(i, ws, W, pos, mask)
            if collect is not None:
                collect.append(ws.xs.float().view(R, T, D).clone())
        P = p * p * m.out_channels
        if out is None:
            out = torch.empty((R, T, P) if train else (R, m.out_channels, H, Wd), dtype=torch.float32, device=x.device)
        sh, sc = ws.fin_views
        with rng("fit/final_layer"):
            if getattr(ws, "fin_mod", None) is not None and W.fin_w16 is not None and P % 8 == 0:
                if ws.fin_c is None or ws.fin_c.shape[1] != P:
                    ws.fin_c = torch.empty(2 * R, P, dtype=torch.float32, device=x.device)
                    ws.fin_tok = torch.empty(R * T, P, dtype=torch.float32, device=x.device)
                # row sums of ws.xs: written by the last block's fc2 epilogue (or row_stats when the network has no blocks)
                ops.modulation_pack(sh, sc, W.n_mod, R, D, 1, 0, ws.fin_mod)
                ops.gemm(ws.fin_mod[0], W.fin_w16, W.fin_b, ws.fin_c, EPI_BIAS, bias_row0=R)
                tok = out.view(R * T, P) if train else ws.fin_tok
                ops.gemm(ws.xs, W.fin_w16, None, tok, EPI_BIAS, rows_per_sample=T, ln=(ws.fin_mod[0, :R], ws.fin_c, ws.parts, LN_EPS))
                if not train:
                    ops.unpatchify(ws.fin_tok.view(R, T, P), out, p)
            else:
                ops.final_layer(ws.xs, sh, sc, W.n_mod, W.fin_w, W.fin_b, out, R, T, p, m.out_channels)
        return out
