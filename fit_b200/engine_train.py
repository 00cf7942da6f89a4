"""Training-step executor: FiT forward that keeps what the backward needs, and the hand-sequenced backward.

Autograd of the reference's ``model(x_t, t, y, pos, mask)`` under ``loss.backward()`` (FiT/train.py:55-66,
masked_FiT/train.py:50-60) as explicit C-ABI launches: every GEMM of the backward is the tcgen05 kernel (dgrad reads
the nn.Linear weight through an MN-major descriptor, wgrad contracts over the token rows with both operands
MN-major), everything else is a kernel of backward.cu / attn_bwd.cu.  No torch arithmetic; torch allocates buffers.

Per block the forward keeps (act dtype = bf16 or fp32):  x_in, LN statistics, h1, qkv (+lse), att, y1 = proj(att),
x_mid, h2, u = packed fc1 pre-activations [M, 2Hp], g, y2 = fc2(g).  FiT-XL/2 at 32 samples x 256 tokens: ~340 MB per
block, 9.5 GB in all -- nothing is recomputed.
"""
from __future__ import annotations

import math
import os
from typing import Callable, Dict, List, Optional

import torch

from . import ops
from ._lib import EPI_BIAS, EPI_BIAS_ROPE
from ._nvtx import rng

Tensor = torch.Tensor


def _round_up(a: int, b: int) -> int:
    return (a + b - 1) // b * b


class TrainWeights:
    """Parameters in kernel layout for the training path (act dtype; fc1_g | fc1_x packed in groups of 16 rows).

    The buffers are allocated ONCE and refilled in place (``refresh``), so their addresses can be handed to the optimizer:
    ``FusedAdamWEMA.attach_engine`` makes the fused AdamW kernel write the bf16 / packed copy of every updated parameter straight into
    them (``shadow_specs``), and nothing re-casts or re-packs the 675 M parameters after a step.  fp32 tensors the kernels read in the
    parameter's own layout (biases of qkv / proj / fc2, the embedders) are views of the parameters themselves."""

    def __init__(self, model, precision: str):
        D, L = model.hidden_size, len(model.blocks)
        self.H = H = model.blocks[0].ffn.hidden_features
        self.Hp = Hp = _round_up(H, 16)
        self.D = D
        wdt = torch.bfloat16 if precision == "bf16" else torch.float32
        self.wdt = wdt
        dev = model.y_embedder.embedding_table.weight.device
        self.t_freqs = torch.exp(-math.log(10000.0) * torch.arange(128, dtype=torch.float32) / 128).to(dev)  # fit.py:73-75
        self.n_mod = 6 * D * L + 2 * D
        self.ada_w = torch.empty(self.n_mod, D, device=dev, dtype=wdt)
        self.ada_b = torch.empty(self.n_mod, device=dev, dtype=torch.float32)
        self.blocks = []
        for _ in model.blocks:
            e = {"qkv_w": torch.empty(3 * D, D, device=dev, dtype=wdt), "proj_w": torch.empty(D, D, device=dev, dtype=wdt),
                 "fc1_w": torch.zeros(2 * Hp, D, device=dev, dtype=wdt), "fc1_b": torch.zeros(2 * Hp, device=dev, dtype=torch.float32),
                 "fc2_w": torch.zeros(D, Hp, device=dev, dtype=wdt)}   # padded rows / columns (FiT-L) stay zero
            self.blocks.append(e)
        self.fin_w = torch.empty(model.final_layer.linear.weight.shape, device=dev, dtype=wdt)
        self.refresh(model)

    def refresh(self, model):
        """(Re)fill every buffer from the fp32 parameters: after construction, load_state_dict, an EMA swap -- any parameter change that did
        not come from the attached optimizer."""
        D, H, Hp = self.D, self.H, self.Hp

        def f32(p):
            return p.detach().float().contiguous()   # fp32 contiguous parameters: a view, not a copy

        self.x_w, self.x_b = f32(model.x_embedder.weight), f32(model.x_embedder.bias)
        self.t0_w, self.t0_b = f32(model.t_embedder.mlp[0].weight), f32(model.t_embedder.mlp[0].bias)
        self.t2_w, self.t2_b = f32(model.t_embedder.mlp[2].weight), f32(model.t_embedder.mlp[2].bias)
        self.y_table = f32(model.y_embedder.embedding_table.weight)
        for i, lin in enumerate(self._ada_linears(model)):
            o, n = self._ada_slice(i)
            self.ada_w[o:o + n].copy_(lin.weight.detach())
            self.ada_b[o:o + n].copy_(lin.bias.detach())
        for blk, e in zip(model.blocks, self.blocks):
            e["qkv_w"].copy_(blk.attn.qkv.weight.detach())
            e["proj_w"].copy_(blk.attn.proj.weight.detach())
            e["qkv_b"], e["proj_b"], e["fc2_b"] = f32(blk.attn.qkv.bias), f32(blk.attn.proj.bias), f32(blk.ffn.fc2.bias)
            w1 = e["fc1_w"].view(Hp // 16, 2, 16, D)
            b1 = e["fc1_b"].view(Hp // 16, 2, 16)
            for half, lin in enumerate((blk.ffn.fc1_g, blk.ffn.fc1_x)):
                full, rem = H // 16, H % 16
                w1[:full, half].copy_(lin.weight.detach()[:full * 16].view(full, 16, D))
                b1[:full, half].copy_(lin.bias.detach()[:full * 16].view(full, 16))
                if rem:
                    w1[full, half, :rem].copy_(lin.weight.detach()[full * 16:])
                    b1[full, half, :rem].copy_(lin.bias.detach()[full * 16:])
            e["fc2_w"][:, :H].copy_(blk.ffn.fc2.weight.detach())
        self.fin_w.copy_(model.final_layer.linear.weight.detach())
        self.fin_b = f32(model.final_layer.linear.bias)

    @staticmethod
    def _ada_linears(model):
        return [b.adaLN_modulation[1] for b in model.blocks] + [model.final_layer.adaLN_modulation[1]]

    def _ada_slice(self, i):
        L = len(self.blocks)
        return (6 * self.D * i, 6 * self.D) if i < L else (6 * self.D * L, 2 * self.D)

    def shadow_specs(self, model):
        """[(parameter, shadow tensor, row_len, shadow_ld, pack_off)]: where the fused optimizer kernel writes the kernel-layout copy of each
        parameter that is not read in place (fitb200_adam_tensor.shadow)."""
        D, Hp = self.D, self.Hp
        specs = []
        for i, lin in enumerate(self._ada_linears(model)):
            o, n = self._ada_slice(i)
            specs.append((lin.weight, self.ada_w[o:o + n], D, D, -1))
            specs.append((lin.bias, self.ada_b[o:o + n], 1, 1, -1))
        for blk, e in zip(model.blocks, self.blocks):
            specs.append((blk.attn.qkv.weight, e["qkv_w"], D, D, -1))
            specs.append((blk.attn.proj.weight, e["proj_w"], D, D, -1))
            specs.append((blk.ffn.fc1_g.weight, e["fc1_w"], D, D, 0))
            specs.append((blk.ffn.fc1_x.weight, e["fc1_w"], D, D, 16))
            specs.append((blk.ffn.fc1_g.bias, e["fc1_b"], 1, 1, 0))
            specs.append((blk.ffn.fc1_x.bias, e["fc1_b"], 1, 1, 16))
            specs.append((blk.ffn.fc2.weight, e["fc2_w"], self.H, Hp, -1))
        specs.append((model.final_layer.linear.weight, self.fin_w, D, D, -1))
        return specs


class Saved:
    """Everything one forward keeps for its backward."""
    pass


class FiTTrainEngine:
    def __init__(self, model, precision: str):
        self.model, self.precision = model, precision
        self._w: Optional[TrainWeights] = None
        self._w_key = None

    @property
    def act(self):
        return torch.bfloat16 if self.precision == "bf16" else torch.float32

    def _side_stream(self, dev):
        st = getattr(self, "_side", None)
        if st is None or st.device != torch.device(dev):
            st = torch.cuda.Stream(device=dev)
            self._side = st
        return st

    def _key(self):
        return (self.precision,) + tuple((p.data_ptr(), p._version) for p in self.model.parameters())

    def weights(self) -> TrainWeights:
        key = self._key()
        if self._w is None:
            if not next(self.model.parameters()).is_cuda:
                raise RuntimeError("fit_b200: the model must live on a CUDA device (no CPU fallback)")
            self._w = TrainWeights(self.model, self.precision)
        elif key != self._w_key:
            self._w.refresh(self.model)     # same buffers, new contents
        self._w_key = key
        return self._w

    def adopt_parameter_versions(self):
        """Called by the attached optimizer after a step whose kernel wrote the kernel-layout copies itself: the buffers are current."""
        if self._w is not None:
            self._w_key = self._key()

    # ------------------------------------------------------------------------------------------ forward
    def forward(self, x: Tensor, t: Tensor, y: Tensor, pos: Tensor, mask: Tensor, train: bool = True):
        m = self.model
        p, D, L = m.patch_size, m.hidden_size, len(m.blocks)
        h, d = m.num_heads, m.hidden_size // m.num_heads
        act, dev = self.act, x.device
        bf16 = self.precision == "bf16"
        if not train:   # latents in: patchify is a view/copy (no arithmetic); the token path below is the same
            Hh, Ww = x.shape[2], x.shape[3]
            x = m.patchify(x)
        x = x.float().contiguous()
        R, T = x.shape[0], x.shape[1]
        M = R * T
        if pos.dim() == 2:
            pos = pos[None]
        pos = pos.float().contiguous()
        if pos.shape[0] not in (1, R) or pos.shape[1] != T or pos.shape[2] != d:
            raise ValueError(f"pos shape {tuple(pos.shape)} does not match (1|{R}, {T}, {d})")
        if tuple(mask.shape) != (R, T):
            raise ValueError(f"mask shape {tuple(mask.shape)} does not match ({R}, {T})")
        W = self.weights()
        nm, Hp = W.n_mod, W.Hp
        S = Saved()
        S.R, S.T, S.pos, S.mask, S.train = R, T, pos, mask, train
        S.hw = None if train else (Hh, Ww)
        t = t.to(torch.int64)
        S.y = y.to(torch.int64).contiguous()
        S.tokens = x.view(M, -1)
        f32 = dict(dtype=torch.float32, device=dev)
        a = dict(dtype=act, device=dev)
        S.kv_len = torch.empty(R, dtype=torch.int32, device=dev)
        ops.mask_to_kvlen(mask, S.kv_len, torch.empty(R, dtype=torch.int32, device=dev))
        xs = torch.empty(M, D, **a)
        ops.embed_tokens(x, W.x_w, W.x_b, xs, R, T, p, m.in_channels)
        # conditioning (fit.py:503-505) with the pre-activation of the t-embedder MLP kept for its SiLU backward
        S.tfreq = torch.empty(R, 256, **f32)
        ops.timestep_sincos(t, S.tfreq, W.t_freqs)
        S.t_pre = torch.empty(R, D, **f32)
        ops.gemm(S.tfreq, W.t0_w, W.t0_b, S.t_pre, EPI_BIAS)
        S.t1 = ops.silu_fwd(S.t_pre, torch.empty(R, D, **f32))
        t2 = torch.empty(R, D, **f32)
        ops.gemm(S.t1, W.t2_w, W.t2_b, t2, EPI_BIAS)
        S.c = torch.empty(R, D, **f32)
        S.silu_c = torch.empty(R, D, **a)
        ops.cond_combine(t2, S.y, W.y_table, S.c, S.silu_c)
        S.mod = torch.empty(R, nm, **f32)
        ops.gemm(S.silu_c, W.ada_w, W.ada_b, S.mod, EPI_BIAS)
        S.blocks = []
        for i in range(L):
          with rng(f"train/fwd/block/{i}"):
            e = W.blocks[i]
            sh1, sc1, g1, sh2, sc2, g2 = [S.mod[:, (6 * i + k) * D:(6 * i + k + 1) * D] for k in range(6)]
            b = Saved()
            b.x_in = xs
            b.mean1, b.rstd1 = torch.empty(M, **f32), torch.empty(M, **f32)
            b.h1 = ops.ln_modulate(xs, sh1, sc1, nm, T, torch.empty(M, D, **a), b.mean1, b.rstd1)
            b.qkv = torch.empty(M, 3 * D, **a)
            if bf16:
                ops.gemm(b.h1, e["qkv_w"], e["qkv_b"], b.qkv, EPI_BIAS_ROPE, rope=pos, rope_cols=2 * D)
            else:
                ops.gemm(b.h1, e["qkv_w"], e["qkv_b"], b.qkv, EPI_BIAS)
            b.att = torch.empty(M, D, **a)
            b.lse = torch.empty(R, h, T, **f32)
            ops.attention(b.qkv, None if bf16 else pos, S.kv_len, mask, R, T, h, d, b.att, b.lse, prerotated=bf16)
            b.y1 = ops.gemm(b.att, e["proj_w"], e["proj_b"], torch.empty(M, D, **a), EPI_BIAS)
            b.x_mid = ops.gate_resid_fwd(xs, b.y1, g1, nm, T, torch.empty(M, D, **a))
            b.mean2, b.rstd2 = torch.empty(M, **f32), torch.empty(M, **f32)
            b.h2 = ops.ln_modulate(b.x_mid, sh2, sc2, nm, T, torch.empty(M, D, **a), b.mean2, b.rstd2)
            b.u = ops.gemm(b.h2, e["fc1_w"], e["fc1_b"], torch.empty(M, 2 * Hp, **a), EPI_BIAS)
            b.g = ops.swiglu_fwd(b.u, torch.empty(M, Hp, **a))
            b.y2 = ops.gemm(b.g, e["fc2_w"], e["fc2_b"], torch.empty(M, D, **a), EPI_BIAS)
            xs = ops.gate_resid_fwd(b.x_mid, b.y2, g2, nm, T, torch.empty(M, D, **a))
            S.blocks.append(b)
        S.x_last = xs
        fsh, fsc = [S.mod[:, 6 * L * D + k * D: 6 * L * D + (k + 1) * D] for k in range(2)]
        S.meanf, S.rstdf = torch.empty(M, **f32), torch.empty(M, **f32)
        S.hf = ops.ln_modulate(xs, fsh, fsc, nm, T, torch.empty(M, D, **a), S.meanf, S.rstdf)
        P = p * p * m.out_channels
        out = torch.empty(M, P, **f32)
        ops.gemm(S.hf, W.fin_w, W.fin_b, out, EPI_BIAS)
        out = out.view(R, T, P)
        if not train:
            out = m.unpatchify(out, Hh, Ww)
        return out, S

    # ------------------------------------------------------------------------------------------ backward
    def backward(self, S: Saved, dout: Tensor, on_grads: Optional[Callable[[str, Dict[str, Tensor]], None]] = None) -> Dict[str, Tensor]:
        """dout: gradient of the forward's output (fp32 or act dtype).  Returns {parameter name: fp32 gradient}.
        on_grads(tag, grads) is called as soon as a group of parameter gradients is final ("final_layer", "blocks.27", ...,
        "blocks.0", "embedders") -- the hook the data-parallel all-reduce overlaps on."""
        m = self.model
        p, D, L = m.patch_size, m.hidden_size, len(m.blocks)
        h, d = m.num_heads, D // m.num_heads
        R, T = S.R, S.T
        M = R * T
        W = self.weights()
        nm, Hp, H = W.n_mod, W.Hp, W.H
        act, dev = self.act, S.x_last.device
        bf16 = self.precision == "bf16"
        f32 = dict(dtype=torch.float32, device=dev)
        a = dict(dtype=act, device=dev)
        grads: Dict[str, Tensor] = {}

        def emit(tag, names, flat=None):
            if on_grads is not None:
                if flat is not None:
                    on_grads(tag, {n: grads[n] for n in names}, flat)
                else:
                    on_grads(tag, {n: grads[n] for n in names})

        def wgrad(dy: Tensor, x: Tensor, out: Optional[Tensor] = None) -> Tensor:           # dW[N, K] = dy[M, N]^T x[M, K]
            return ops.gemm(dy, x, None, torch.empty(dy.shape[1], x.shape[1], **f32) if out is None else out, EPI_BIAS, a_mn=True, b_mn=True)

        # The weight-gradient GEMMs of a block run on a SIDE stream next to the dgrad chain: at 8192 token rows the persistent GEMM grids end in
        # partial waves (proj / fc2 dgrad: 160 tiles on 74 CTA pairs = 2.16 waves; dW_proj: 25 tiles = a third of the GPU), and a second kernel's
        # CTAs start on the SMs the first one has left.  Nothing downstream reads a weight gradient before the block's buckets are handed over,
        # so the side stream is joined once per block.  Tensors it reads are kept alive until that join (`keep`): the caching allocator may
        # hand a freed block to the main stream only after the join has ordered the two streams.  FITB200_TRAIN_WGRAD_STREAM=0 switches it off.
        main_stream = torch.cuda.current_stream(dev)
        side = self._side_stream(dev) if os.environ.get("FITB200_TRAIN_WGRAD_STREAM", "1") != "0" else None
        keep: List[Tensor] = []

        class _Side:
            def __enter__(_s):
                if side is not None:
                    side.wait_stream(main_stream)          # everything queued so far (the dy this wgrad reads) is ordered before it
                    _s.ctx = torch.cuda.stream(side)
                    _s.ctx.__enter__()

            def __exit__(_s, *a):
                if side is not None:
                    _s.ctx.__exit__(*a)

        def join_side():
            if side is not None:
                main_stream.wait_stream(side)
            keep.clear()

        def bucket(shapes):
            """One flat fp32 buffer holding the listed gradients back to back (16-byte aligned offsets): the data-parallel all-reduce runs
            on it in place -- no packing copy before and no scatter copy after the collective.  Only the vectors (bias gradients, accumulated
            with atomics) and the alignment gaps are zeroed: the weight gradients are overwritten whole by the wgrad GEMMs (zero-filling
            the 96 MB bucket of every block was 2.7 GB of memset per step)."""
            offs, n = [], 0
            for shp in shapes:
                offs.append(n)
                cnt = 1
                for v in shp:
                    cnt *= v
                n += (cnt + 3) // 4 * 4
            flat = torch.empty(n, **f32)
            views = []
            for off, shp in zip(offs, shapes):
                cnt = 1
                for v in shp:
                    cnt *= v
                if len(shp) == 1:
                    flat[off:off + (cnt + 3) // 4 * 4].zero_()
                views.append(flat[off:off + cnt].view(*shp))
            return flat, views

        def dgrad(dy: Tensor, w: Tensor, out_dtype=None) -> Tensor:  # dx[M, K] = dy[M, N] w[N, K]
            return ops.gemm(dy, w, None, torch.empty(dy.shape[0], w.shape[1], dtype=out_dtype or act, device=dev), EPI_BIAS, b_mn=True)

        P = p * p * m.out_channels
        if not S.train:
            dout = m.patchify(dout)
        dout = dout.reshape(M, P).contiguous()
        if dout.dtype != act:
            dout = ops.cast(dout.float().contiguous() if dout.dtype not in (torch.float32, torch.bfloat16) else dout, torch.empty(M, P, **a))
        dmod = torch.zeros(R, nm, **f32)

        # ---- final layer (dit.py:138-142)
        grads["final_layer.linear.weight"] = wgrad(dout, S.hf)
        grads["final_layer.linear.bias"] = ops.colsum(dout, torch.zeros(P, **f32))
        dhf = dgrad(dout, W.fin_w)
        o = 6 * L * D
        dx = ops.ln_modulate_bwd(dhf, S.x_last, S.meanf, S.rstdf, S.mod[:, o + D:o + 2 * D], nm, T, None, torch.empty(M, D, **a),
                                 dmod[:, o:o + D], dmod[:, o + D:o + 2 * D], nm)
        emit("final_layer", ["final_layer.linear.weight", "final_layer.linear.bias"])

        # ---- blocks, last to first (fit.py:326-337)
        for i in reversed(range(L)):
          with rng(f"train/bwd/block/{i}"):
            e, b = W.blocks[i], S.blocks[i]
            pre = f"blocks.{i}."
            mv = [S.mod[:, (6 * i + k) * D:(6 * i + k + 1) * D] for k in range(6)]
            dv = [dmod[:, (6 * i + k) * D:(6 * i + k + 1) * D] for k in range(6)]
            flat, (g_fc2b, g_fc2w, g_f1gw, g_f1xw, g_f1gb, g_f1xb, g_pb, g_pw, g_qb, g_qw, g_aw, g_ab) = bucket(
                [(D,), (D, H), (H, D), (H, D), (H,), (H,), (D,), (D, D), (3 * D,), (3 * D, D), (6 * D, D), (6 * D,)])
            # x_out = x_mid + gate_mlp * fc2(g)
            dy2 = ops.gate_bwd(dx, b.y2, mv[5], nm, R, T, torch.empty(M, D, **a), dv[5], nm, g_fc2b)
            grads[pre + "ffn.fc2.bias"] = g_fc2b
            with _Side():
                if Hp != H:
                    tmp = wgrad(dy2, b.g)
                    g_fc2w.copy_(tmp[:, :H])
                    keep.append(tmp)
                else:
                    wgrad(dy2, b.g, g_fc2w)
            keep.extend([dy2, b.g])
            grads[pre + "ffn.fc2.weight"] = g_fc2w
            dg = dgrad(dy2, e["fc2_w"])
            # g = silu(u_gate) * u_up
            db1 = torch.zeros(2 * Hp, **f32)
            du = ops.swiglu_bwd(dg, b.u, torch.empty(M, 2 * Hp, **a), db1)
            with _Side():
                dw1 = wgrad(du, b.h2).view(Hp // 16, 2, 16, D)
                g_f1gw.copy_(dw1[:, 0].reshape(Hp, D)[:H]); g_f1xw.copy_(dw1[:, 1].reshape(Hp, D)[:H])   # un-interleave the packed fc1 rows
            keep.extend([du, b.h2, dw1])
            grads[pre + "ffn.fc1_g.weight"], grads[pre + "ffn.fc1_x.weight"] = g_f1gw, g_f1xw
            db1 = db1.view(Hp // 16, 2, 16)
            g_f1gb.copy_(db1[:, 0].reshape(Hp)[:H]); g_f1xb.copy_(db1[:, 1].reshape(Hp)[:H])
            grads[pre + "ffn.fc1_g.bias"], grads[pre + "ffn.fc1_x.bias"] = g_f1gb, g_f1xb
            dh2 = dgrad(du, e["fc1_w"])
            dx_mid = ops.ln_modulate_bwd(dh2, b.x_mid, b.mean2, b.rstd2, mv[4], nm, T, dx, torch.empty(M, D, **a), dv[3], dv[4], nm)
            # x_mid = x_in + gate_msa * proj(att)
            dy1 = ops.gate_bwd(dx_mid, b.y1, mv[2], nm, R, T, torch.empty(M, D, **a), dv[2], nm, g_pb)
            grads[pre + "attn.proj.bias"] = g_pb
            with _Side():
                grads[pre + "attn.proj.weight"] = wgrad(dy1, b.att, g_pw)
            keep.extend([dy1, b.att])
            datt = dgrad(dy1, e["proj_w"])
            dqkv = ops.attention_bwd(b.qkv, S.pos, S.kv_len, S.mask, R, T, h, d, b.att, datt, b.lse, torch.empty(M, 3 * D, **a), prerotated=bf16)
            grads[pre + "attn.qkv.bias"] = ops.colsum(dqkv, g_qb)
            with _Side():
                grads[pre + "attn.qkv.weight"] = wgrad(dqkv, b.h1, g_qw)
            keep.extend([dqkv, b.h1])
            dh1 = dgrad(dqkv, e["qkv_w"])
            dx = ops.ln_modulate_bwd(dh1, b.x_in, b.mean1, b.rstd1, mv[1], nm, T, dx_mid, torch.empty(M, D, **a), dv[0], dv[1], nm)
            # adaLN table of this block: mod_i = silu(c) W_i^T + b_i (fit.py:323-324,329).  Its six dmod chunks are final now, so the
            # gradient (a third of all parameters) joins this block's bucket instead of waiting for the end of the backward
            dmod_i = dmod[:, 6 * i * D:6 * (i + 1) * D]
            dmod_ia = dmod_i.contiguous() if act == torch.float32 else ops.cast(dmod_i.contiguous(), torch.empty(R, 6 * D, **a))
            grads[pre + "adaLN_modulation.1.weight"] = wgrad(dmod_ia, S.silu_c, g_aw)
            grads[pre + "adaLN_modulation.1.bias"] = ops.colsum(dmod_i, g_ab)
            join_side()
            emit(pre[:-1], [pre + s for s in ("ffn.fc2.bias", "ffn.fc2.weight", "ffn.fc1_g.weight", "ffn.fc1_x.weight", "ffn.fc1_g.bias",
                                              "ffn.fc1_x.bias", "attn.proj.bias", "attn.proj.weight", "attn.qkv.bias", "attn.qkv.weight",
                                              "adaLN_modulation.1.weight", "adaLN_modulation.1.bias")], flat)
            S.blocks[i] = None  # free this block's activations

        # ---- x_embedder (fit.py:387-388,499-501): no gradient with respect to the latents is needed
        tok = S.tokens if S.tokens.dtype == act else ops.cast(S.tokens, torch.empty(S.tokens.shape, **a))
        grads["x_embedder.weight"] = wgrad(dx, tok)
        grads["x_embedder.bias"] = ops.colsum(dx, torch.zeros(D, **f32))

        # ---- adaLN tables of all blocks + final layer: mod = silu(c) W_ada^T + b  (fit.py:323-324,329; dit.py:134-139)
        dmod_a = dmod if act == torch.float32 else ops.cast(dmod, torch.empty(R, nm, **a))
        dmod_f = dmod[:, 6 * L * D:].contiguous()
        dmod_fa = dmod_f if act == torch.float32 else ops.cast(dmod_f, torch.empty(R, 2 * D, **a))
        grads["final_layer.adaLN_modulation.1.weight"] = wgrad(dmod_fa, S.silu_c)
        grads["final_layer.adaLN_modulation.1.bias"] = ops.colsum(dmod_f, torch.zeros(2 * D, **f32))
        dsilu_c = dgrad(dmod_a, W.ada_w, torch.float32)
        dc = ops.silu_bwd(dsilu_c, S.c, torch.empty(R, D, **f32))
        # ---- c = t_emb + y_emb (fit.py:505): label table scatter, t-embedder MLP (fit.py:55-57,84-85)
        grads["y_embedder.embedding_table.weight"] = ops.embedding_bwd(dc, S.y, torch.zeros(W.y_table.shape, **f32))
        grads["t_embedder.mlp.2.bias"] = ops.colsum(dc, torch.zeros(D, **f32))
        grads["t_embedder.mlp.2.weight"] = wgrad(dc, S.t1)
        dt1 = dgrad(dc, W.t2_w, torch.float32)
        dpre = ops.silu_bwd(dt1, S.t_pre, torch.empty(R, D, **f32))
        grads["t_embedder.mlp.0.bias"] = ops.colsum(dpre, torch.zeros(D, **f32))
        grads["t_embedder.mlp.0.weight"] = wgrad(dpre, S.tfreq)
        emit("embedders", [n for n in grads if not n.startswith("blocks.") and not n.startswith("final_layer.linear")])
        return grads
