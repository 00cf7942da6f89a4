"""Tensor-level wrappers over the C ABI: argument checking, pointers, the current CUDA stream.

Each function enqueues one kernel of libfit_b200.so on ``torch.cuda.current_stream()``; the caller
allocates outputs (these helpers allocate only when ``out`` is None).  No torch arithmetic here.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib as L
from ._lib import BF16, EPI_BIAS, EPI_BIAS_GATE_RESID, EPI_BIAS_ROPE, EPI_BIAS_SILU, EPI_SWIGLU, F32  # noqa: F401

Tensor = torch.Tensor


def _c(t: Tensor) -> Tensor:
    if not t.is_contiguous():
        raise ValueError("fit_b200: tensor must be contiguous")
    return t


def mask_to_kvlen(mask: Tensor, kv_len: Optional[Tensor] = None, n_valid: Optional[Tensor] = None):
    L.require_cuda(mask)
    R, T = mask.shape
    m8 = _c(mask.view(torch.uint8) if mask.dtype == torch.bool else mask)
    if kv_len is None:
        kv_len = torch.empty(R, dtype=torch.int32, device=mask.device)
    if n_valid is None:
        n_valid = torch.empty(R, dtype=torch.int32, device=mask.device)
    L.check(L.lib().fitb200_mask_to_kvlen(L.ptr(m8), R, T, L.ptr(kv_len), L.ptr(n_valid), L.stream_ptr()), "mask_to_kvlen")
    return kv_len, n_valid


def embed_tokens(x: Tensor, w: Tensor, b: Tensor, out: Tensor, R: int, T: int, p: int, C_in: int):
    """x: tokens [R_src,T,p*p*C] or latents [R_src,C,H,W] (fp32) -> out [R,T,D] (row r reads r % R_src)."""
    L.require_cuda(x, w, b, out)
    nchw = x.dim() == 4
    R_src = x.shape[0]
    H, W = (x.shape[2], x.shape[3]) if nchw else (0, 0)
    D = w.shape[0]
    L.check(L.lib().fitb200_embed_tokens_fwd(L.ptr(_c(x)), int(nchw), R, R_src, C_in, H, W, p, T, L.ptr(_c(w)), L.ptr(_c(b)),
                                             D, L.ptr(out), L.dtype_code(out), L.stream_ptr()), "embed_tokens")
    return out


def timestep_sincos(t: Tensor, out: Tensor, freqs: Optional[Tensor] = None):
    L.require_cuda(t, out)
    assert t.dtype == torch.int64 and out.dtype == torch.float32
    L.check(L.lib().fitb200_timestep_sincos(L.ptr(_c(t)), t.shape[0], L.ptr(freqs), L.ptr(out), L.stream_ptr()), "timestep_sincos")
    return out


def cond_combine(t_emb: Tensor, y: Tensor, table: Tensor, c: Optional[Tensor], silu_c: Tensor):
    L.require_cuda(t_emb, y, table, silu_c)
    R, D = silu_c.shape
    assert y.dtype == torch.int64 and t_emb.dtype == torch.float32 and table.dtype == torch.float32
    L.check(L.lib().fitb200_cond_combine_fwd(L.ptr(_c(t_emb)), t_emb.shape[0], L.ptr(_c(y)), L.ptr(_c(table)), table.shape[0],
                                             R, D, L.ptr(c), L.ptr(silu_c), L.dtype_code(silu_c), L.stream_ptr()), "cond_combine")
    return silu_c


def gemm(A: Tensor, B: Tensor, bias: Optional[Tensor], out: Tensor, epilogue: int = EPI_BIAS, *, M: Optional[int] = None,
         B2: Optional[Tensor] = None, bias2: Optional[Tensor] = None, gate: Optional[Tensor] = None, ld_gate: int = 0,
         rows_per_sample: int = 1, resid: Optional[Tensor] = None, N: Optional[int] = None,
         rope: Optional[Tensor] = None, rope_cols: int = 0, a_mn: bool = False, b_mn: bool = False, K: Optional[int] = None,
         bias_row0: int = 0, ln: Optional[tuple] = None, stats_out: Optional[Tensor] = None):
    """out = epilogue(A @ B^T).  A [M,K] and B [N,K] both fp32 (CUDA-core path) or both bf16 (tcgen05 path).
    ``gate`` may be a strided view: pass the storage-offset view and its row stride ``ld_gate``.
    ``ln`` = (g, c, parts, eps): LayerNorm + adaLN-modulate prologue -- A is the raw residual stream, g bf16 [S, K] = 1 + scale,
    c fp32 [2S, N] the (c1, c2) vectors, parts fp32 [M, nparts, 2] the partial row sums (fitb200_gemm_args.ln_*); needs rows_per_sample.
    ``stats_out`` (GATE_RESID): fp32 [M, gemm_stats_parts(N), 2] partial row sums of the output."""
    L.require_cuda(A, B, out)
    a = L.GemmArgs()
    # a_mn / b_mn: the operand is stored contraction-major, [K, M] resp. [K, N] (backward GEMMs)
    a.M = (A.shape[1] if a_mn else A.shape[0]) if M is None else M
    a.K = (A.shape[0] if a_mn else A.shape[1]) if K is None else K
    a.N = (B.shape[1] if b_mn else B.shape[0]) if N is None else N
    a.a_mn, a.b_mn = int(a_mn), int(b_mn)
    a.A, a.lda = A.data_ptr(), A.stride(0)
    a.B, a.ldb = B.data_ptr(), B.stride(0)
    a.B2 = 0 if B2 is None else B2.data_ptr()
    a.bias = 0 if bias is None else bias.data_ptr()
    a.bias2 = 0 if bias2 is None else bias2.data_ptr()
    a.out, a.ldo, a.out_dtype = out.data_ptr(), out.stride(0), L.dtype_code(out)
    a.epilogue = epilogue
    a.gate = 0 if gate is None else gate.data_ptr()
    a.ld_gate, a.rows_per_sample = ld_gate, rows_per_sample
    a.resid = 0 if resid is None else resid.data_ptr()
    a.ld_resid = 0 if resid is None else resid.stride(0)
    a.resid_dtype = F32 if resid is None else L.dtype_code(resid)
    if rope is not None:   # EPI_BIAS_ROPE: fp32 table [1|samples, tokens, head_dim]
        assert rope.dtype == torch.float32 and rope.dim() == 3 and rope.is_contiguous()
        a.rope_freqs, a.rope_rows, a.rope_tokens, a.rope_head_dim, a.rope_cols = rope.data_ptr(), rope.shape[0], rope.shape[1], rope.shape[2], rope_cols
    a.bias_row0 = bias_row0
    if ln is not None:
        g, c, parts, eps = ln
        assert g.dtype == torch.bfloat16 and g.dim() == 2 and g.stride(1) == 1 and g.shape[1] == a.K
        assert c.dtype == torch.float32 and c.is_contiguous() and c.shape == (2 * g.shape[0], a.N)
        assert parts.dtype == torch.float32 and parts.is_contiguous() and parts.dim() == 3 and parts.shape[0] >= a.M and parts.shape[2] == 2
        a.ln_g, a.ld_ln_g = g.data_ptr(), g.stride(0)
        a.ln_c, a.ln_samples = c.data_ptr(), g.shape[0]
        a.ln_parts, a.ln_nparts, a.ln_eps = parts.data_ptr(), parts.shape[1], eps
    if stats_out is not None:
        assert stats_out.dtype == torch.float32 and stats_out.is_contiguous() and stats_out.shape == (a.M, gemm_stats_parts(a.N), 2)
        a.stats_out = stats_out.data_ptr()
    if A.dtype == torch.float32 and B.dtype == torch.float32:
        fn, what = L.lib().fitb200_gemm_f32, "gemm_f32"
    elif A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16:
        fn, what = L.lib().fitb200_gemm_bf16_tc, "gemm_bf16_tc"
    else:
        raise TypeError(f"fit_b200.gemm: A {A.dtype} / B {B.dtype} must both be fp32 or both bf16")
    L.check(fn(C.byref(a), L.stream_ptr()), what)
    return out


def unpatchify(tokens: Tensor, out: Tensor, p: int):
    """tokens fp32 [R, T, p*p*C] -> out fp32 [R, C, H, W] (index map only)."""
    L.require_cuda(tokens, out)
    R, C, H, W = out.shape
    assert tokens.dtype == torch.float32 and out.dtype == torch.float32 and tokens.is_contiguous() and out.is_contiguous()
    L.check(L.lib().fitb200_unpatchify(L.ptr(tokens), R, tokens.shape[1], C, H, W, p, L.ptr(out), L.stream_ptr()), "unpatchify")
    return out


def gemm_stats_parts(N: int) -> int:
    """Partial sums per row that a GATE_RESID GEMM with N output columns writes through ``stats_out``."""
    return int(L.lib().fitb200_gemm_stats_parts(N))


def row_stats(x: Tensor, parts: Tensor):
    """parts[m, 0] = (sum x[m], sum x[m]^2), parts[m, 1:] = 0; x [M, D], parts fp32 [M, nparts, 2]."""
    L.require_cuda(x, parts)
    M, D = x.shape
    assert parts.dtype == torch.float32 and parts.is_contiguous() and parts.shape[0] == M and parts.shape[2] == 2
    L.check(L.lib().fitb200_row_stats(L.ptr(_c(x)), L.dtype_code(x), M, D, L.ptr(parts), parts.shape[1], L.stream_ptr()), "row_stats")
    return parts


def modulation_pack(shift: Tensor, scale: Tensor, ld_mod: int, S: int, D: int, n_sets: int, set_stride: int, out: Tensor):
    """out[set] = [bf16(1 + scale_set) ; bf16(shift_set)] ([2S, D] each); shift / scale: fp32 views of the adaLN table (row stride ld_mod)."""
    L.require_cuda(shift, scale, out)
    assert out.dtype == torch.bfloat16 and out.is_contiguous() and out.numel() == n_sets * 2 * S * D
    L.check(L.lib().fitb200_modulation_pack(L.ptr(shift), L.ptr(scale), ld_mod, S, D, n_sets, set_stride, L.ptr(out), L.stream_ptr()), "modulation_pack")
    return out


def ln_modulate(x: Tensor, shift: Tensor, scale: Tensor, ld_mod: int, rows_per_sample: int, out: Tensor,
                mean: Optional[Tensor] = None, rstd: Optional[Tensor] = None):
    """out = LN(x) * (1 + scale[s]) + shift[s]; x, out: [M, D]; shift/scale: fp32 views with row stride ld_mod."""
    L.require_cuda(x, shift, scale, out)
    M, D = x.shape
    L.check(L.lib().fitb200_ln_modulate_fwd(L.ptr(_c(x)), L.dtype_code(x), M, D, rows_per_sample, L.ptr(shift), L.ptr(scale),
                                            ld_mod, L.ptr(out), L.dtype_code(out), L.ptr(mean), L.ptr(rstd), L.stream_ptr()),
            "ln_modulate")
    return out


def ln_apply(x: Tensor, parts: Tensor, shift: Tensor, scale: Tensor, ld_mod: int, rows_per_sample: int, out: Tensor, eps: float = 1e-6):
    """ln_modulate for rows whose partial (sum, sum of squares) are in ``parts`` [M, nparts, 2] (gemm(..., stats_out=parts) / row_stats)."""
    L.require_cuda(x, parts, shift, scale, out)
    M, D = x.shape
    assert parts.dtype == torch.float32 and parts.is_contiguous() and parts.shape[0] == M and parts.shape[2] == 2
    L.check(L.lib().fitb200_ln_apply_fwd(L.ptr(_c(x)), L.dtype_code(x), M, D, rows_per_sample, L.ptr(parts), parts.shape[1], eps, L.ptr(shift),
                                         L.ptr(scale), ld_mod, L.ptr(out), L.dtype_code(out), L.stream_ptr()), "ln_apply")
    return out


def attention(qkv: Tensor, freqs: Optional[Tensor], kv_len: Tensor, mask: Optional[Tensor], R: int, T: int, h: int, d: int,
              out: Tensor, lse: Optional[Tensor] = None, prerotated: bool = False):
    """qkv [R*T, 3*h*d] -> out [R*T, h*d]; RoPE + key-masked softmax attention.
    prerotated: q / k already rotated by a gemm(..., EPI_BIAS_ROPE) (bf16 path)."""
    L.require_cuda(qkv, kv_len, out)
    assert kv_len.dtype == torch.int32
    if prerotated:
        L.check(L.lib().fitb200_attn_fwd(L.ptr(_c(qkv)), L.dtype_code(qkv), None, 1, L.ptr(kv_len),
                                         L.ptr(None if mask is None else _c(mask.view(torch.uint8) if mask.dtype == torch.bool else mask)),
                                         R, T, h, d, 1, L.ptr(out), L.ptr(lse), L.stream_ptr()), "attn_fwd")
        return out
    L.require_cuda(freqs)
    assert freqs.dtype == torch.float32
    m8 = None
    if mask is not None:
        m8 = _c(mask.view(torch.uint8) if mask.dtype == torch.bool else mask)
    fr = 1 if freqs.dim() == 2 else freqs.shape[0]
    L.check(L.lib().fitb200_attn_fwd(L.ptr(_c(qkv)), L.dtype_code(qkv), L.ptr(_c(freqs)), fr, L.ptr(kv_len), L.ptr(m8), R, T, h, d, 0,
                                     L.ptr(out), L.ptr(lse), L.stream_ptr()), "attn_fwd")
    return out


def final_layer(x: Tensor, shift: Tensor, scale: Tensor, ld_mod: int, w: Tensor, b: Tensor, out: Tensor, R: int, T: int,
                p: int, C_out: int):
    """x [R*T, D] -> out tokens [R,T,P] or latents [R,C,H,W] (decided by out.dim())."""
    L.require_cuda(x, shift, scale, w, b, out)
    D = x.shape[-1]
    P = w.shape[0]
    nchw = out.dim() == 4
    H, W = (out.shape[2], out.shape[3]) if nchw else (0, 0)
    assert out.dtype == torch.float32 and w.dtype == torch.float32
    L.check(L.lib().fitb200_final_layer_fwd(L.ptr(_c(x)), L.dtype_code(x), R, T, D, L.ptr(shift), L.ptr(scale), ld_mod, L.ptr(_c(w)),
                                            L.ptr(_c(b)), P, L.ptr(out), int(nchw), C_out, H, W, p, L.stream_ptr()), "final_layer")
    return out


STEP_CLIP_DENOISED, STEP_PREDICT_XSTART = 1, 2


def cfg_step(x: Tensor, model_out: Tensor, noise: Optional[Tensor], t: Tensor, coef: Tensor, *, cfg_pairs: int = 0,
             cfg_scale: float = 1.0, ddim: bool = False, eta: float = 0.0, flags: int = 0, sample: Optional[Tensor] = None,
             pred_xstart: Optional[Tensor] = None, mean: Optional[Tensor] = None, want_sample: bool = True):
    """Fused (CFG combine +) DDPM / DDIM update.  Returns (sample, pred_xstart, mean)."""
    L.require_cuda(x, model_out, t, coef)
    assert x.dtype == torch.float32 and model_out.dtype == torch.float32
    assert noise is None or noise.dtype == torch.float32
    assert t.dtype == torch.int64 and coef.dtype == torch.float32 and coef.shape[1] == 8
    rows = x.shape[0]
    per_row = x.numel() // rows
    if sample is None and want_sample:
        sample = torch.empty_like(x)
    S = coef.shape[0]
    noise = None if noise is None else _c(noise)
    if ddim:
        rc = L.lib().fitb200_cfg_ddim_step(L.ptr(_c(x)), L.ptr(_c(model_out)), L.ptr(noise), L.ptr(sample), L.ptr(pred_xstart),
                                           rows, per_row, cfg_pairs, cfg_scale, eta, flags, L.ptr(_c(t)), L.ptr(_c(coef)), S,
                                           L.stream_ptr())
    else:
        rc = L.lib().fitb200_cfg_ddpm_step(L.ptr(_c(x)), L.ptr(_c(model_out)), L.ptr(noise), L.ptr(sample), L.ptr(pred_xstart),
                                           L.ptr(mean), rows, per_row, cfg_pairs, cfg_scale, flags, L.ptr(_c(t)),
                                           L.ptr(_c(coef)), S, L.stream_ptr())
    L.check(rc, "cfg_step")
    return sample, pred_xstart, mean


def cfg_combine(model_out: Tensor, cfg_scale: float, out: Optional[Tensor] = None):
    """forward_with_cfg's eps combine: rows [0,B) cond, [B,2B) uncond -> both halves = u + s (c - u)."""
    L.require_cuda(model_out)
    pairs = model_out.shape[0] // 2
    if out is None:
        out = torch.empty_like(model_out)
    L.check(L.lib().fitb200_cfg_combine(L.ptr(_c(model_out)), L.ptr(out), pairs, model_out.numel() // model_out.shape[0],
                                        float(cfg_scale), L.stream_ptr()), "cfg_combine")
    return out


def q_sample(x0: Tensor, noise: Tensor, t: Tensor, sqrt_ac: Tensor, sqrt_1m_ac: Tensor, out: Optional[Tensor] = None):
    L.require_cuda(x0, noise, t, sqrt_ac, sqrt_1m_ac)
    rows = x0.shape[0]
    if out is None:
        out = torch.empty_like(x0)
    L.check(L.lib().fitb200_q_sample(L.ptr(_c(x0)), L.ptr(_c(noise)), L.ptr(out), rows, x0.numel() // rows, L.ptr(_c(t)),
                                     L.ptr(sqrt_ac), L.ptr(sqrt_1m_ac), sqrt_ac.shape[0], L.stream_ptr()), "q_sample")
    return out


def philox_normal(out: Tensor, seed: int, step: Optional[Tensor], row_offset=0):
    """out[rows, ...] <- N(0,1) keyed by (seed, step[0], global row index).  row_offset: int, or a 1-element int64 device tensor
    (read at run time: a captured graph then serves any batch position)."""
    L.require_cuda(out)
    assert out.dtype == torch.float32 and out.is_contiguous()
    rows = out.shape[0]
    off_dev = None
    if isinstance(row_offset, torch.Tensor):
        assert row_offset.dtype == torch.int64 and row_offset.is_cuda
        off_dev, row_offset = row_offset, 0
    L.check(L.lib().fitb200_philox_normal(L.ptr(out), rows, out.numel() // rows, seed & (2 ** 64 - 1), L.ptr(step), int(row_offset),
                                          L.ptr(off_dev), L.stream_ptr()), "philox_normal")
    return out


def gather_i64(table: Tensor, idx: Tensor, out: Tensor):
    L.check(L.lib().fitb200_gather_i64(L.ptr(table), table.shape[0], L.ptr(idx), L.ptr(out), idx.shape[0], L.stream_ptr()), "gather_i64")
    return out


def add_i64(idx: Tensor, delta: int):
    L.check(L.lib().fitb200_add_i64(L.ptr(idx), idx.shape[0], int(delta), L.stream_ptr()), "add_i64")
    return idx


# ---------------------------------------------------------------------------------------------- training step (backward.cu, attn_bwd.cu)
def _m8(mask):
    return None if mask is None else _c(mask.view(torch.uint8) if mask.dtype == torch.bool else mask)


def attention_bwd(qkv: Tensor, freqs: Tensor, kv_len: Tensor, mask: Optional[Tensor], R: int, T: int, h: int, d: int, out: Tensor,
                  dout: Tensor, lse: Tensor, dqkv: Tensor, prerotated: bool = False, force_simt: bool = False):
    """Gradient of attention() with respect to the un-rotated q, k and v: dqkv [R*T, 3*h*d]."""
    L.require_cuda(qkv, freqs, kv_len, out, dout, lse, dqkv)
    assert freqs.dtype == torch.float32 and lse.dtype == torch.float32 and qkv.dtype == out.dtype == dout.dtype == dqkv.dtype
    fr = 1 if freqs.dim() == 2 else freqs.shape[0]
    L.check(L.lib().fitb200_attn_bwd(L.ptr(_c(qkv)), L.dtype_code(qkv), L.ptr(_c(freqs)), fr, L.ptr(kv_len), L.ptr(_m8(mask)), R, T, h, d,
                                     int(prerotated) | (2 if force_simt else 0), L.ptr(_c(out)), L.ptr(_c(dout)), L.ptr(lse), L.ptr(dqkv),
                                     L.stream_ptr()), "attn_bwd")
    return dqkv


def gate_resid_fwd(resid: Tensor, y: Tensor, gate: Tensor, ld_gate: int, rows_per_sample: int, out: Tensor):
    M, D = resid.shape
    L.check(L.lib().fitb200_gate_resid_fwd(L.ptr(_c(resid)), L.ptr(_c(y)), L.dtype_code(resid), L.ptr(gate), ld_gate, M, D, rows_per_sample,
                                           L.ptr(out), L.stream_ptr()), "gate_resid_fwd")
    return out


def gate_bwd(dx: Tensor, y: Tensor, gate: Tensor, ld_gate: int, R: int, rows_per_sample: int, dy: Tensor, dgate: Tensor, ld_dgate: int,
             dbias: Optional[Tensor]):
    D = dx.shape[1]
    L.check(L.lib().fitb200_gate_bwd(L.ptr(_c(dx)), L.ptr(_c(y)), L.dtype_code(dx), L.ptr(gate), ld_gate, R, rows_per_sample, D, L.ptr(dy),
                                     L.ptr(dgate), ld_dgate, L.ptr(dbias), L.stream_ptr()), "gate_bwd")
    return dy


def swiglu_fwd(u: Tensor, g: Tensor):
    M, Hp = g.shape
    L.check(L.lib().fitb200_swiglu_fwd(L.ptr(_c(u)), L.dtype_code(u), M, Hp, L.ptr(g), L.stream_ptr()), "swiglu_fwd")
    return g


def swiglu_bwd(dg: Tensor, u: Tensor, du: Tensor, dbias: Optional[Tensor]):
    M, Hp = dg.shape
    L.check(L.lib().fitb200_swiglu_bwd(L.ptr(_c(dg)), L.ptr(_c(u)), L.dtype_code(u), M, Hp, L.ptr(du), L.ptr(dbias), L.stream_ptr()), "swiglu_bwd")
    return du


def ln_modulate_bwd(dh: Tensor, x: Tensor, mean: Tensor, rstd: Tensor, scale: Tensor, ld_mod: int, rows_per_sample: int,
                    dres: Optional[Tensor], dx: Tensor, dshift: Optional[Tensor], dscale: Optional[Tensor], ld_dmod: int):
    M, D = x.shape
    L.check(L.lib().fitb200_ln_modulate_bwd(L.ptr(_c(dh)), L.ptr(_c(x)), L.dtype_code(x), L.ptr(mean), L.ptr(rstd), L.ptr(scale), ld_mod, M, D,
                                            rows_per_sample, L.ptr(dres), L.ptr(dx), L.ptr(dshift), L.ptr(dscale), ld_dmod, L.stream_ptr()),
            "ln_modulate_bwd")
    return dx


def colsum(a: Tensor, out: Tensor):
    """out[n] += sum_m a[m, n] (out fp32, zero-initialised by the caller)."""
    M, N = a.shape
    L.check(L.lib().fitb200_colsum(L.ptr(a), L.dtype_code(a), M, N, a.stride(0), L.ptr(out), L.stream_ptr()), "colsum")
    return out


def silu_fwd(x: Tensor, y: Tensor):
    L.check(L.lib().fitb200_silu_fwd(L.ptr(_c(x)), x.numel(), L.ptr(y), L.stream_ptr()), "silu_fwd")
    return y


def silu_bwd(dy: Tensor, x: Tensor, dx: Tensor):
    L.check(L.lib().fitb200_silu_bwd(L.ptr(_c(dy)), L.ptr(_c(x)), x.numel(), L.ptr(dx), L.stream_ptr()), "silu_bwd")
    return dx


def embedding_bwd(dc: Tensor, y: Tensor, dtable: Tensor):
    R, D = dc.shape
    L.check(L.lib().fitb200_embedding_bwd(L.ptr(_c(dc)), L.ptr(_c(y)), R, D, dtable.shape[0], L.ptr(dtable), L.stream_ptr()), "embedding_bwd")
    return dtable


def cast(src: Tensor, dst: Tensor):
    L.check(L.lib().fitb200_cast(L.ptr(_c(src)), L.dtype_code(src), L.ptr(dst), L.dtype_code(dst), src.numel(), L.stream_ptr()), "cast")
    return dst


def masked_mse_fwd(out: Tensor, target: Tensor, mask: Tensor, acc: Tensor, loss: Tensor):
    P = out.shape[-1]
    L.check(L.lib().fitb200_masked_mse_fwd(L.ptr(_c(out)), L.ptr(_c(target)), L.ptr(_m8(mask)), mask.numel(), P, L.ptr(acc), L.ptr(loss),
                                           L.stream_ptr()), "masked_mse_fwd")
    return loss


def masked_mse_bwd(out: Tensor, target: Tensor, mask: Tensor, acc: Tensor, gscale: Optional[Tensor], dout: Tensor):
    P = out.shape[-1]
    L.check(L.lib().fitb200_masked_mse_bwd(L.ptr(_c(out)), L.ptr(_c(target)), L.ptr(_m8(mask)), mask.numel(), P, L.ptr(acc), L.ptr(gscale),
                                           L.ptr(dout), L.dtype_code(dout), L.stream_ptr()), "masked_mse_bwd")
    return dout


def subsample_tokens(latent: Tensor, pos: Tensor, mask: Tensor, perm: Tensor, n_tokens: Tensor, number_of_tokens: int):
    """masked_FiT subsample + collate on device (masked_FiT/preprocess/iterators.py:142-169): latent [N,Tsrc,P] fp32, pos [N,Tsrc,d] fp32,
    mask [N,Tsrc] bool, perm [N,Tsrc] int64 (host-drawn randperm of range(n_tokens[s]), unused where n_tokens[s] <= T'), n_tokens [N] int32.
    Returns (latent [N,T',P], pos [N,T',d], mask [N,T'] bool)."""
    L.require_cuda(latent, pos, mask, perm, n_tokens)
    N, Tsrc, P = latent.shape
    d = pos.shape[2]
    assert pos.shape[:2] == (N, Tsrc) and mask.shape == (N, Tsrc) and perm.shape == (N, Tsrc) and n_tokens.shape == (N,)
    assert latent.dtype == torch.float32 and pos.dtype == torch.float32 and perm.dtype == torch.int64 and n_tokens.dtype == torch.int32
    lo = torch.empty(N, number_of_tokens, P, dtype=torch.float32, device=latent.device)
    po = torch.empty(N, number_of_tokens, d, dtype=torch.float32, device=latent.device)
    mo = torch.empty(N, number_of_tokens, dtype=torch.bool, device=latent.device)
    L.check(L.lib().fitb200_subsample_tokens(L.ptr(_c(latent)), L.ptr(_c(pos)), L.ptr(_m8(mask)), L.ptr(_c(perm)), L.ptr(n_tokens), N, Tsrc, P, d,
                                             number_of_tokens, L.ptr(lo), L.ptr(po), L.ptr(mo), L.stream_ptr()), "subsample_tokens")
    return lo, po, mo
