"""GPU parity of the individual C-ABI kernels against the CPU oracle (run on the B200 box: -m gpu)."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from helpers import golden, rel_l2
from oracle import diffusion_oracle as do
from oracle import fit_oracle as fo

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from fit_b200 import ops as _ops
    from fit_b200 import _lib
    assert _lib.lib().fitb200_device_ok() == 1, "not a compute-capability 10.x device"
    return _ops


def _coef(diff, device="cuda"):
    return diff._tables(torch.device(device))


# ------------------------------------------------------------------------------ diffusion step (family 4)
def test_step_kernels_match_reference_golden(ops):
    from fit_b200.diffusion import create_diffusion
    z = golden("diffusion_steps")
    d = create_diffusion("10")
    x, eps, noise = (torch.tensor(z[k]).cuda() for k in ("x", "eps", "noise"))
    t = torch.tensor(z["t"]).cuda()
    model = lambda x_, t_, **kw: eps  # noqa: E731
    o = d.p_sample(model, x, t, clip_denoised=False, noise=noise)
    # fp32 elementwise: tolerance 1e-6 relative; exp() differs from CPU libm by <= 2 ulp
    assert rel_l2(o["sample"], z["p_sample.sample"]) < 1e-6
    assert rel_l2(o["pred_xstart"], z["p_sample.pred_xstart"]) < 1e-6
    np.testing.assert_array_equal(o["pred_xstart"].cpu().numpy(), z["p_sample.pred_xstart"])  # no transcendental: bit-exact
    pm = d.p_mean_variance(model, x, t, clip_denoised=False)
    np.testing.assert_array_equal(pm["mean"].cpu().numpy(), z["pmv.mean"])
    np.testing.assert_array_equal(pm["variance"].cpu().numpy(), z["pmv.variance"])
    np.testing.assert_array_equal(pm["log_variance"].cpu().numpy(), z["pmv.log_variance"])
    for eta in (0.0, 0.5):
        o = d.ddim_sample(model, x, t, clip_denoised=False, eta=eta, noise=noise)
        assert rel_l2(o["sample"], z[f"ddim{eta}.sample"]) < 1e-6
    np.testing.assert_array_equal(d.q_sample(x, t, noise=noise).cpu().numpy(), z["q_sample"])


@pytest.mark.parametrize("rows,per_row", [(2, 4), (8, 4 * 32 * 32), (6, 4 * 20 * 40), (512, 4 * 32 * 32)])
def test_fused_cfg_step_vs_oracle(ops, rows, per_row):
    from fit_b200.diffusion import create_diffusion
    d = create_diffusion("250")
    tb, _ = do.create("250")
    g = torch.Generator().manual_seed(rows)
    B = rows // 2
    x = torch.randn(rows, per_row, generator=g)
    raw = torch.randn(rows, per_row, generator=g)
    noise = torch.randn(rows, per_row, generator=g)
    t = torch.randint(0, 250, (rows,), generator=g)
    t[0] = 0
    cond, unc = raw[:B], raw[B:]
    eps = unc + 1.5 * (cond - unc)
    eps = torch.cat([eps, eps], 0)
    want = do.p_sample(tb, x, t, eps, noise)
    coef = _coef(d)["coef"]
    s, x0, _ = ops.cfg_step(x.cuda(), raw.cuda(), noise.cuda(), t.cuda(), coef, cfg_pairs=B, cfg_scale=1.5,
                            pred_xstart=torch.empty(rows, per_row, device="cuda"))
    assert rel_l2(s, want["sample"]) < 1e-6
    np.testing.assert_array_equal(x0.cpu().numpy(), want["pred_xstart"].numpy())
    want = do.ddim_sample(tb, x, t, eps, noise, eta=0.3)
    s, _, _ = ops.cfg_step(x.cuda(), raw.cuda(), noise.cuda(), t.cuda(), coef, cfg_pairs=B, cfg_scale=1.5, ddim=True, eta=0.3)
    assert rel_l2(s, want["sample"]) < 1e-6
    # in-place update (sample aliases x), as the graph-captured sampler uses it
    xc = x.cuda().clone()
    ops.cfg_step(xc, raw.cuda(), noise.cuda(), t.cuda(), coef, cfg_pairs=B, cfg_scale=1.5, sample=xc)
    assert rel_l2(xc, do.p_sample(tb, x, t, eps, noise)["sample"]) < 1e-6
    out = ops.cfg_combine(raw.cuda().contiguous(), 1.5)
    np.testing.assert_array_equal(out.cpu().numpy(), eps.numpy())


def test_step_flags_clip_and_xstart(ops):
    from fit_b200.diffusion import create_diffusion
    d = create_diffusion("10", predict_xstart=True)
    tb, _ = do.create("10")
    g = torch.Generator().manual_seed(3)
    x, mo, noise = (torch.randn(4, 64, generator=g) * 2 for _ in range(3))
    t = torch.tensor([9, 3, 0, 5])
    o = d.p_sample(lambda x_, t_, **kw: mo.cuda(), x.cuda(), t.cuda(), clip_denoised=True, noise=noise.cuda())
    x0 = mo.clamp(-1, 1)
    mean = do._gather(tb.posterior_mean_coef1, t, 2) * x0 + do._gather(tb.posterior_mean_coef2, t, 2) * x
    nz = (t != 0).float()[:, None]
    want = mean + nz * torch.exp(0.5 * do._gather(tb.fixed_large_log_variance, t, 2)) * noise
    assert rel_l2(o["sample"], want) < 1e-6
    np.testing.assert_array_equal(o["pred_xstart"].cpu().numpy(), x0.numpy())


def test_timestep_map_gather(ops):
    from fit_b200.diffusion import create_diffusion
    d = create_diffusion("250")
    t = torch.tensor([0, 1, 249, 100], device="cuda")
    np.testing.assert_array_equal(d._model_timesteps(t).cpu().numpy(), np.array(d.timestep_map)[[0, 1, 249, 100]])
    idx = torch.tensor([5, 5, 5], device="cuda")
    ops.add_i64(idx, -1)
    assert idx.tolist() == [4, 4, 4]


# ------------------------------------------------------------------------------ elementwise (family 3)
def test_mask_to_kvlen(ops):
    m = torch.zeros(5, 300, dtype=torch.bool)
    m[0, :256] = True
    m[1, :1] = True
    m[3, ::7] = True
    m[4] = True
    kv, nv = ops.mask_to_kvlen(m.cuda())
    assert kv.tolist() == [256, 1, 0, 295, 300]
    assert nv.tolist() == [256, 1, 0, int(m[3].sum()), 300]


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("p,C,H,W,D", [(2, 4, 32, 32, 1152), (2, 4, 20, 40, 768), (4, 4, 16, 16, 384), (2, 4, 10, 6, 144)])
def test_embed_tokens(ops, dt, p, C, H, W, D):
    g = torch.Generator().manual_seed(D + H)
    x = torch.randn(3, C, H, W, generator=g)
    w = torch.randn(D, p * p * C, generator=g) * 0.3
    b = torch.randn(D, generator=g)
    T = (H // p) * (W // p)
    want = F.linear(fo.patchify(x, p), w, b)
    tol = 1e-6 if dt == torch.float32 else 4e-3
    out = torch.empty(3 * T, D, dtype=dt, device="cuda")
    ops.embed_tokens(x.cuda(), w.cuda(), b.cuda(), out, 3, T, p, C)
    assert rel_l2(out.float().view(3, T, D), want) < tol
    # token input + row duplication (forward_with_cfg's cat([half, half]))
    out2 = torch.empty(6 * T, D, dtype=dt, device="cuda")
    ops.embed_tokens(fo.patchify(x, p).contiguous().cuda(), w.cuda(), b.cuda(), out2, 6, T, p, C)
    assert rel_l2(out2.float().view(6, T, D), torch.cat([want, want], 0)) < tol


def test_timestep_and_conditioning(ops):
    t = torch.tensor([0, 1, 17, 500, 999, 333])
    out = torch.empty(6, 256, device="cuda")
    want = fo.timestep_embedding(t)
    # with the reference's host-computed frequency table (fit.py:73-75) the arguments are bit-identical and
    # fp32 sin/cos of CUDA vs CPU libm agree to ~1e-6 absolute for arguments up to 999 rad
    freqs = torch.exp(-math.log(10000.0) * torch.arange(128, dtype=torch.float32) / 128)
    ops.timestep_sincos(t.cuda(), out, freqs.cuda())
    assert (out.cpu() - want).abs().max() < 2e-6
    # in-kernel frequencies: expf differs from the host's by <= 1 ulp, times t <= 999 rad
    ops.timestep_sincos(t.cuda(), out)
    assert (out.cpu() - want).abs().max() < 2e-4
    g = torch.Generator().manual_seed(0)
    D = 144
    temb = torch.randn(6, D, generator=g)
    table = torch.randn(11, D, generator=g)
    y = torch.tensor([0, 10, 3, 3, 9, 1])
    c = torch.empty(6, D, device="cuda")
    sc = torch.empty(6, D, device="cuda")
    ops.cond_combine(temb.cuda(), y.cuda(), table.cuda(), c, sc)
    np.testing.assert_array_equal(c.cpu().numpy(), (temb + table[y]).numpy())
    assert rel_l2(sc, F.silu(temb + table[y])) < 1e-6
    # batch-uniform timestep (1 row of t_emb broadcast)
    ops.cond_combine(temb[:1].contiguous().cuda(), y.cuda(), table.cuda(), c, sc)
    np.testing.assert_array_equal(c.cpu().numpy(), (temb[:1] + table[y]).numpy())


@pytest.mark.parametrize("xdt,odt", [(torch.float32, torch.float32), (torch.bfloat16, torch.bfloat16), (torch.float32, torch.bfloat16)])
@pytest.mark.parametrize("D,T", [(1152, 256), (768, 200), (128, 64), (144, 17)])
def test_ln_modulate(ops, xdt, odt, D, T):
    g = torch.Generator().manual_seed(D)
    R = 3
    x = (torch.randn(R, T, D, generator=g) * 3 + 0.5).to(xdt)
    mod = torch.randn(R, 6 * D, generator=g) * 0.5
    want = fo.modulate(fo.layer_norm(x.float()), mod[:, :D], mod[:, D:2 * D])
    modc = mod.cuda()
    out = torch.empty(R * T, D, dtype=odt, device="cuda")
    mean = torch.empty(R * T, device="cuda")
    rstd = torch.empty(R * T, device="cuda")
    ops.ln_modulate(x.cuda().view(R * T, D), modc[:, :D], modc[:, D:2 * D], 6 * D, T, out, mean, rstd)
    tol = 2e-6 if odt == torch.float32 else 4e-3
    assert rel_l2(out.float().view(R, T, D), want) < tol
    assert rel_l2(mean, x.float().mean(-1).reshape(-1)) < 1e-5


@pytest.mark.parametrize("xdt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("D,H,W,p", [(1152, 32, 32, 2), (768, 20, 40, 2), (144, 8, 8, 4)])
def test_final_layer(ops, xdt, D, H, W, p):
    g = torch.Generator().manual_seed(D + 1)
    R, C = 2, 4
    T = (H // p) * (W // p)
    P = p * p * C
    x = (torch.randn(R, T, D, generator=g) * 2).to(xdt)
    mod = torch.randn(R, 2 * D, generator=g) * 0.5
    w = torch.randn(P, D, generator=g) * 0.05
    b = torch.randn(P, generator=g)
    tok = F.linear(fo.modulate(fo.layer_norm(x.float()), mod[:, :D], mod[:, D:]), w, b)
    modc = mod.cuda()
    out = torch.empty(R, T, P, device="cuda")
    ops.final_layer(x.cuda().view(R * T, D), modc[:, :D], modc[:, D:], 2 * D, w.cuda(), b.cuda(), out, R, T, p, C)
    assert rel_l2(out, tok) < 1e-5
    out4 = torch.empty(R, C, H, W, device="cuda")
    ops.final_layer(x.cuda().view(R * T, D), modc[:, :D], modc[:, D:], 2 * D, w.cuda(), b.cuda(), out4, R, T, p, C)
    assert rel_l2(out4, fo.unpatchify(tok, H, W, p, C)) < 1e-5


@pytest.mark.parametrize("xdt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("R,D,H,W,p", [(17, 1152, 32, 32, 2), (33, 768, 20, 40, 2), (24, 384, 32, 32, 2), (17, 1024, 64, 64, 4)])
def test_final_layer_persistent_smem_weight(ops, xdt, R, D, H, W, p):
    """>= 4096 tokens: the persistent kernel with the [P, D] weight in shared memory (two tokens per warp and round; odd token counts, every
    FiT width, P = 16 and 64) -- same arithmetic as the one-warp-per-token kernel, which the small shapes above take."""
    g = torch.Generator().manual_seed(D + R)
    C = 4
    T = (H // p) * (W // p)
    P = p * p * C
    assert R * T >= 4096
    x = (torch.randn(R, T, D, generator=g) * 2).to(xdt)
    mod = torch.randn(R, 2 * D, generator=g) * 0.5
    w = torch.randn(P, D, generator=g) * 0.05
    b = torch.randn(P, generator=g)
    tok = F.linear(fo.modulate(fo.layer_norm(x.float()), mod[:, :D], mod[:, D:]), w, b)
    modc = mod.cuda()
    out = torch.empty(R, T, P, device="cuda")
    ops.final_layer(x.cuda().view(R * T, D), modc[:, :D], modc[:, D:], 2 * D, w.cuda(), b.cuda(), out, R, T, p, C)
    assert rel_l2(out, tok) < 1e-5
    out4 = torch.empty(R, C, H, W, device="cuda")
    ops.final_layer(x.cuda().view(R * T, D), modc[:, :D], modc[:, D:], 2 * D, w.cuda(), b.cuda(), out4, R, T, p, C)
    assert rel_l2(out4, fo.unpatchify(tok, H, W, p, C)) < 1e-5


# ------------------------------------------------------------------------------ attention (family 2)
def _attn_case(R, T, h, d, lens, seed, nonprefix=False):
    g = torch.Generator().manual_seed(seed)
    qkv = torch.randn(R, T, 3, h, d, generator=g)
    from oracle.pos_embed_oracle import precompute_freqs_cis_2d
    side = int(np.ceil(np.sqrt(T)))
    tab = torch.tensor(precompute_freqs_cis_2d(d, side, side, max_length=256), dtype=torch.float32)[:T]
    freqs = torch.zeros(R, T, d)
    mask = torch.zeros(R, T, dtype=torch.bool)
    for r, L in enumerate(lens):
        freqs[r, :L] = tab[:L]
        mask[r, :L] = True
    if nonprefix:
        mask[0, 1:T:3] = False
    return qkv, freqs, mask


def _attn_oracle(qkv, freqs, mask):
    q, k, v = qkv.permute(2, 0, 3, 1, 4)
    q, k = fo.rope_rotate(q, freqs), fo.rope_rotate(k, freqs)
    o = fo.attention_core(q, k, v, mask)
    return o.transpose(1, 2).reshape(qkv.shape[0], qkv.shape[1], -1)


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("T,h,d,lens,nonprefix", [
    (256, 16, 72, [256, 200, 1, 130], False),
    (256, 12, 64, [256, 200, 256, 65], False),
    (64, 2, 64, [64, 50, 64, 36], True),
    (400, 4, 72, [400, 400], False),
    (96, 3, 72, [96, 10, 33], True),
    (256, 2, 72, [16, 17, 128, 129, 144, 255], False),
    (200, 2, 64, [200, 128, 7, 160], True),
    (8, 1, 64, [8, 3], False),
])
def test_attention(ops, dt, T, h, d, lens, nonprefix):
    R = len(lens)
    qkv, freqs, mask = _attn_case(R, T, h, d, lens, seed=T + d, nonprefix=nonprefix)
    if dt == torch.bfloat16:
        qkv = qkv.bfloat16().float()   # same rounded inputs on both sides
    kv, _ = ops.mask_to_kvlen(mask.cuda())
    out = torch.empty(R * T, h * d, dtype=dt, device="cuda")
    lse = torch.empty(R, h, T, device="cuda")
    q3, k3, v3 = qkv.permute(2, 0, 3, 1, 4)
    if dt == torch.float32:
        # verification mode: un-rotated q / k + the (cos, sin) table, RoPE applied in the kernel
        want = _attn_oracle(qkv, freqs, mask)
        qr, kr = fo.rope_rotate(q3, freqs), fo.rope_rotate(k3, freqs)
        ops.attention(qkv.cuda().view(R * T, 3 * h * d), freqs.cuda(), kv, mask.cuda(), R, T, h, d, out, lse)
    else:
        # bf16: the operands carry the RoPE already (the qkv GEMM's EPI_BIAS_ROPE epilogue wrote rope(q), rope(k)): TMA-fed tcgen05 kernels
        qr, kr = fo.rope_rotate(q3, freqs).bfloat16().float(), fo.rope_rotate(k3, freqs).bfloat16().float()
        want = fo.attention_core(qr, kr, v3, mask).transpose(1, 2).reshape(R, T, -1)
        qkv_rot = torch.stack([qr, kr, v3], 0).permute(1, 3, 0, 2, 4).contiguous().bfloat16()   # [R, T, 3, h, d]
        ops.attention(qkv_rot.cuda().view(R * T, 3 * h * d), None, kv, mask.cuda(), R, T, h, d, out, lse, prerotated=True)
        with pytest.raises(RuntimeError, match="carry the RoPE"):   # un-rotated bf16 operands: refused, not silently mis-computed
            ops.attention(qkv.to(dt).cuda().view(R * T, 3 * h * d), freqs.cuda(), kv, mask.cuda(), R, T, h, d, torch.empty_like(out))
    out = out.float().cpu().view(R, T, h * d)
    tol = 2e-5 if dt == torch.float32 else 5e-3
    for r, L in enumerate(lens):   # compare the valid tokens only
        assert rel_l2(out[r, :L], want[r, :L]) < tol, (r, L)
    # log-sum-exp of the scaled, key-masked scores (saved for the backward pass)
    sc = (qr @ kr.transpose(-1, -2)) * d ** -0.5
    sc = sc.masked_fill(~mask[:, None, None, :], float("-inf"))
    want_lse = torch.logsumexp(sc, -1)
    for r, L in enumerate(lens):
        assert (lse[r, :, :L].cpu() - want_lse[r, :, :L]).abs().max() < (1e-4 if dt == torch.float32 else 3e-2), (r, L)
    # broadcast table (freqs_rows = 1) on full-length rows (fp32: the table is an input of the kernel)
    if dt == torch.float32 and all(L == T for L in lens) and not nonprefix:
        out1 = torch.empty(R * T, h * d, dtype=dt, device="cuda")
        ops.attention(qkv.cuda().view(R * T, 3 * h * d), freqs[:1].contiguous().cuda(), kv, None, R, T, h, d, out1)
        assert rel_l2(out1.float().cpu().view(R, T, h * d), want) < tol


@pytest.mark.parametrize("h,d", [(16, 72), (12, 64)])
def test_attention_persistent_many_items(ops, h, d):
    """attn_tc2 (persistent: several (sample, head) items per CTA): ring-buffer reuse, alternating softmax groups and skipped items with
    ragged key counts -- 0 (item skipped), <= 128 (one query unit) and > 128 (two units) mixed so that unit parity and ring slots drift."""
    T, R = 256, 40
    g = torch.Generator().manual_seed(17)
    lens = [256, 0, 1, 128, 129, 200, 16, 0, 255, 64] + [int(x) for x in torch.randint(0, 257, (R - 10,), generator=g)]
    qkv, freqs, mask = _attn_case(R, T, h, d, lens, seed=3, nonprefix=False)
    q3, k3, v3 = qkv.bfloat16().float().permute(2, 0, 3, 1, 4)
    qr, kr = fo.rope_rotate(q3, freqs).bfloat16().float(), fo.rope_rotate(k3, freqs).bfloat16().float()
    want = fo.attention_core(qr, kr, v3, mask).transpose(1, 2).reshape(R, T, -1)
    qkv_rot = torch.stack([qr, kr, v3], 0).permute(1, 3, 0, 2, 4).contiguous().bfloat16()
    kv, _ = ops.mask_to_kvlen(mask.cuda())
    out = torch.full((R * T, h * d), float("nan"), dtype=torch.bfloat16, device="cuda")
    lse = torch.full((R, h, T), float("nan"), device="cuda")
    ops.attention(qkv_rot.cuda().view(R * T, 3 * h * d), None, kv, mask.cuda(), R, T, h, d, out, lse, prerotated=True)
    assert bool(torch.isfinite(out.float()).all()) and bool(torch.isfinite(lse).all())   # every row written, padded ones included
    out = out.float().cpu().view(R, T, h * d)
    for r, L in enumerate(lens):
        if L > 0:
            assert rel_l2(out[r, :L], want[r, :L]) < 5e-3, (r, L)
        q0 = (L + 127) // 128 * 128
        assert float(out[r, q0:].abs().max()) == 0.0 if q0 < T else True   # query tiles beyond kv_len are zero rows


@pytest.mark.parametrize("T,h,d,nonprefix", [(400, 16, 72, False), (1000, 3, 72, True), (640, 4, 64, False), (300, 2, 64, True)])
def test_attention_kv_blocked_long(ops, T, h, d, nonprefix):
    """attn_tc3 (T > 256, pre-rotated q / k: the extrapolation shapes of BASELINE config 5): online softmax over 128-key blocks with
    O rescaled in TMEM, ragged key counts around every block boundary (0 = item skipped, 1, 128, 129, 256, 257, T - 1, T), a
    non-prefix mask, ring / phase drift over many units per CTA, and the saved log-sum-exp."""
    g = torch.Generator().manual_seed(T + d)
    base = [T, 0, 1, 128, 129, 256, 257, T - 1, 17, 385 if T > 385 else T - 3]
    lens = base + [int(x) for x in torch.randint(0, T + 1, (14,), generator=g)]
    R = len(lens)
    qkv, freqs, mask = _attn_case(R, T, h, d, lens, seed=5, nonprefix=nonprefix)
    # scores with a spread large enough that the running maximum keeps moving between blocks
    qkv[:, :, 0] *= 2.0
    q3, k3, v3 = qkv.bfloat16().float().permute(2, 0, 3, 1, 4)
    qr, kr = fo.rope_rotate(q3, freqs).bfloat16().float(), fo.rope_rotate(k3, freqs).bfloat16().float()
    want = fo.attention_core(qr, kr, v3, mask).transpose(1, 2).reshape(R, T, -1)
    sc = (qr @ kr.transpose(-1, -2)) * d ** -0.5
    want_lse = torch.logsumexp(sc.masked_fill(~mask[:, None, None, :], float("-inf")), -1)
    qkv_rot = torch.stack([qr, kr, v3], 0).permute(1, 3, 0, 2, 4).contiguous().bfloat16()
    kv, _ = ops.mask_to_kvlen(mask.cuda())
    out = torch.full((R * T, h * d), float("nan"), dtype=torch.bfloat16, device="cuda")
    lse = torch.full((R, h, T), float("nan"), device="cuda")
    ops.attention(qkv_rot.cuda().view(R * T, 3 * h * d), None, kv, mask.cuda(), R, T, h, d, out, lse, prerotated=True)
    assert bool(torch.isfinite(out.float()).all()) and bool(torch.isfinite(lse).all())   # every row written, padded ones included
    out = out.float().cpu().view(R, T, h * d)
    for r, L in enumerate(lens):
        Lm = int(mask[r].sum())
        if Lm > 0:
            assert rel_l2(out[r, :L], want[r, :L]) < 5e-3, (r, L)
            assert (lse[r, :, :L].cpu() - want_lse[r, :, :L]).abs().max() < 3e-2, (r, L)
        q0 = (L + 127) // 128 * 128
        if q0 < T:
            assert float(out[r, q0:].abs().max()) == 0.0   # query tiles beyond kv_len are zero rows


@pytest.mark.parametrize("xdt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("M,D,rps", [(512, 1152, 256), (77, 144, 20), (4100, 768, 200)])
def test_ln_apply_from_row_sums(ops, xdt, M, D, rps):
    """modulate(LayerNorm(x)) from known row sums (row_stats here; the GATE_RESID GEMM epilogue in the engine) == the row-reduction kernel."""
    g = torch.Generator().manual_seed(M + D)
    S = (M + rps - 1) // rps
    x = (torch.randn(M, D, generator=g) * 1.5 + 0.3).to(xdt)
    mod = torch.randn(S, 2 * D, generator=g) * 0.4
    xd = x.double()
    h = F.layer_norm(xd, (D,), eps=1e-6)
    want = h * (1 + mod[:, D:].double().repeat_interleave(rps, 0)[:M]) + mod[:, :D].double().repeat_interleave(rps, 0)[:M]
    modc = mod.cuda()
    parts = torch.full((M, 3, 2), float("nan"), device="cuda")
    ops.row_stats(x.cuda(), parts)
    for odt in (torch.float32, torch.bfloat16):
        out = torch.empty(M, D, dtype=odt, device="cuda")
        ops.ln_apply(x.cuda(), parts, modc[:, :D], modc[:, D:], 2 * D, rps, out)
        assert rel_l2(out.float(), want) < (1e-5 if odt == torch.float32 else 4e-3)
