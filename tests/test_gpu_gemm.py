"""GPU parity of the GEMM kernels (family 1): fp32 CUDA-core verification GEMM and the bf16 tcgen05 GEMM,
every fused epilogue, tile tails in M / N / K.  Floating point: the checker is fp64 math on the same
(bf16-rounded) inputs; tolerance = output rounding (fp32 1e-5, bf16 4e-3 relative L2)."""
import pytest
import torch
import torch.nn.functional as F

from helpers import rel_l2

pytestmark = pytest.mark.gpu

EPI_BIAS, EPI_BIAS_SILU, EPI_BIAS_GATE_RESID, EPI_SWIGLU = 0, 1, 2, 3


@pytest.fixture(scope="module")
def ops():
    from fit_b200 import ops as _ops
    return _ops


def _inputs(M, N, K, seed, wdt):
    g = torch.Generator().manual_seed(seed)
    A = torch.randn(M, K, generator=g).to(wdt)
    B = (torch.randn(N, K, generator=g) / K ** 0.5).to(wdt)
    bias = torch.randn(N, generator=g) * 0.1
    return A, B, bias


SHAPES = [(128, 128, 64), (256, 384, 1152), (200, 3456, 1152), (1000, 1152, 3072), (77, 776, 72), (8, 2304, 768),
          (512, 256, 2736)]


@pytest.mark.parametrize("wdt,odt", [(torch.float32, torch.float32), (torch.bfloat16, torch.bfloat16), (torch.bfloat16, torch.float32)])
@pytest.mark.parametrize("M,N,K", SHAPES)
def test_gemm_bias_and_silu(ops, wdt, odt, M, N, K):
    A, B, bias = _inputs(M, N, K, M + N + K, wdt)
    want = A.double() @ B.double().t() + bias.double()
    tol = 1e-5 if odt == torch.float32 else 4e-3
    out = torch.empty(M, N, dtype=odt, device="cuda")
    ops.gemm(A.cuda(), B.cuda(), bias.cuda(), out, EPI_BIAS)
    assert rel_l2(out.float(), want) < tol
    ops.gemm(A.cuda(), B.cuda(), bias.cuda(), out, EPI_BIAS_SILU)
    assert rel_l2(out.float(), F.silu(want)) < tol


@pytest.mark.parametrize("wdt,adt", [(torch.float32, torch.float32), (torch.bfloat16, torch.bfloat16), (torch.bfloat16, torch.float32)])
@pytest.mark.parametrize("M,N,K,rps", [(512, 1152, 1152, 256), (400, 768, 2048, 200), (128, 128, 64, 64), (96, 144, 384, 32),
                                       (1024, 1152, 3072, 256),
                                       (16640, 1152, 128, 256),     # 325 tiles: several per CTA pair (TMEM stage / ring phases wrap), cut last column tile
                                       (13200, 864, 192, 200)])     # ragged last row block, tiles straddle samples
def test_gemm_gate_residual(ops, wdt, adt, M, N, K, rps):
    A, B, bias = _inputs(M, N, K, M + N, wdt)
    g = torch.Generator().manual_seed(1)
    R = M // rps
    mod = torch.randn(R, 3 * N, generator=g)
    resid = torch.randn(M, N, generator=g).to(adt)
    gate = mod[:, N:2 * N]
    want = resid.double() + gate.double().repeat_interleave(rps, 0) * (A.double() @ B.double().t() + bias.double())
    modc = mod.cuda()
    xs = resid.cuda().clone()
    # in place on the residual stream, gate is a strided view into the adaLN table
    ops.gemm(A.cuda(), B.cuda(), bias.cuda(), xs, EPI_BIAS_GATE_RESID, gate=modc[:, N:2 * N], ld_gate=3 * N,
             rows_per_sample=rps, resid=xs)
    assert rel_l2(xs.float(), want) < (1e-5 if adt == torch.float32 else 4e-3)


@pytest.mark.parametrize("M,K,H", [(256, 1152, 3072), (200, 768, 2048), (130, 128, 341), (64, 1024, 2730), (16484, 128, 1024)])
def test_gemm_swiglu_fp32(ops, M, K, H):
    g = torch.Generator().manual_seed(H)
    A = torch.randn(M, K, generator=g)
    Wg, Wx = torch.randn(H, K, generator=g) / K ** 0.5, torch.randn(H, K, generator=g) / K ** 0.5
    bg, bx = torch.randn(H, generator=g) * 0.1, torch.randn(H, generator=g) * 0.1
    want = F.silu(A.double() @ Wg.double().t() + bg.double()) * (A.double() @ Wx.double().t() + bx.double())
    Hp = (H + 15) // 16 * 16
    out = torch.zeros(M, Hp, device="cuda")
    ops.gemm(A.cuda(), Wg.cuda(), bg.cuda(), out, EPI_SWIGLU, B2=Wx.cuda(), bias2=bx.cuda())
    assert rel_l2(out[:, :H], want) < 1e-5
    assert float(out[:, H:].abs().sum()) == 0.0


@pytest.mark.parametrize("M,K,H", [(256, 1152, 3072), (200, 768, 2048), (130, 128, 341), (64, 1024, 2730), (16484, 128, 1024)])
def test_gemm_swiglu_bf16_packed(ops, M, K, H):
    """Packed [2*Hp, K] weight: 16 rows of fc1_g then the matching 16 rows of fc1_x (the engine's layout)."""
    g = torch.Generator().manual_seed(H + 1)
    A = torch.randn(M, K, generator=g).bfloat16()
    Wg = (torch.randn(H, K, generator=g) / K ** 0.5).bfloat16()
    Wx = (torch.randn(H, K, generator=g) / K ** 0.5).bfloat16()
    bg, bx = torch.randn(H, generator=g) * 0.1, torch.randn(H, generator=g) * 0.1
    want = F.silu(A.double() @ Wg.double().t() + bg.double()) * (A.double() @ Wx.double().t() + bx.double())
    Hp = (H + 15) // 16 * 16
    wg, wx = torch.zeros(Hp, K, dtype=torch.bfloat16), torch.zeros(Hp, K, dtype=torch.bfloat16)
    wg[:H], wx[:H] = Wg, Wx
    b1, b2 = torch.zeros(Hp), torch.zeros(Hp)
    b1[:H], b2[:H] = bg, bx
    Wp = torch.stack([wg.view(Hp // 16, 16, K), wx.view(Hp // 16, 16, K)], 1).reshape(2 * Hp, K).contiguous()
    bp = torch.stack([b1.view(Hp // 16, 16), b2.view(Hp // 16, 16)], 1).reshape(2 * Hp).contiguous()
    out = torch.empty(M, Hp, dtype=torch.bfloat16, device="cuda")
    ops.gemm(A.cuda(), Wp.cuda(), bp.cuda(), out, EPI_SWIGLU)
    assert rel_l2(out[:, :H].float(), want) < 4e-3
    assert float(out[:, H:].float().abs().sum()) == 0.0


def test_gemm_tc_large_persistent(ops):
    """More tiles than SMs (persistent loop, both TMEM accumulator stages, smem ring wrap-around)."""
    M, N, K = 4096, 3456, 1152
    A, B, bias = _inputs(M, N, K, 5, torch.bfloat16)
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    ops.gemm(A.cuda(), B.cuda(), bias.cuda(), out, EPI_BIAS)
    want = (A.cuda().float() @ B.cuda().float().t() + bias.cuda())   # checker at size: fp32 matmul on the same inputs
    assert rel_l2(out.float(), want) < 4e-3
    # linearity (size-independent property): gemm(2A) - bias == 2 (gemm(A) - bias) up to bf16 rounding
    out2 = torch.empty(M, N, dtype=torch.float32, device="cuda")
    out1 = torch.empty(M, N, dtype=torch.float32, device="cuda")
    ops.gemm((A * 2).cuda(), B.cuda(), None, out2, EPI_BIAS)
    ops.gemm(A.cuda(), B.cuda(), None, out1, EPI_BIAS)
    assert torch.equal(out2, out1 * 2)


def test_gemm_argument_errors(ops):
    A = torch.zeros(8, 12, dtype=torch.bfloat16, device="cuda")
    B = torch.zeros(16, 12, dtype=torch.bfloat16, device="cuda")
    out = torch.zeros(8, 16, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(RuntimeError, match="multiples of 8"):
        ops.gemm(A, B, None, out, EPI_BIAS)
    with pytest.raises(TypeError):
        ops.gemm(A.float(), B, None, out, EPI_BIAS)
    with pytest.raises(RuntimeError, match="CUDA tensors"):
        ops.gemm(A.cpu(), B.cpu(), None, out.cpu(), EPI_BIAS)


@pytest.mark.parametrize("odt", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("R,T,h,d,rows", [(2, 256, 16, 72, 1), (3, 200, 12, 64, 3), (2, 64, 2, 72, 2), (4, 33, 2, 64, 1),
                                          (70, 256, 4, 72, 1), (66, 200, 2, 64, 66)])   # many tiles per CTA pair; smem RoPE rows reloaded per row block
def test_gemm_qkv_rope_epilogue(ops, odt, R, T, h, d, rows):
    """EPI_BIAS_ROPE: qkv Linear (fit.py:234) + apply_rotary_emb on the q and k columns (fit.py:160-174), v untouched."""
    from oracle import fit_oracle as fo
    from oracle.pos_embed_oracle import precompute_freqs_cis_2d
    import numpy as np
    D, M = h * d, R * T
    A, B, bias = _inputs(M, 3 * D, D, R * T + d, torch.bfloat16)
    g = torch.Generator().manual_seed(3)
    side = int(np.ceil(np.sqrt(T)))
    tab = torch.tensor(precompute_freqs_cis_2d(d, side, side, max_length=256), dtype=torch.float32)[:T]
    freqs = tab[None].repeat(rows, 1, 1).contiguous()
    if rows > 1:   # per-sample tables (mixed resolution / masked_FiT permuted tokens): make them differ
        for r in range(rows):
            freqs[r] = tab[torch.randperm(T, generator=g)]
    lin = (A.double() @ B.double().t() + bias.double()).view(R, T, 3, h, d)
    q, k, v = lin.permute(2, 0, 3, 1, 4)
    fr = freqs.double()
    want = torch.stack([fo.rope_rotate(q, fr), fo.rope_rotate(k, fr), v], 0).permute(1, 3, 0, 2, 4).reshape(M, 3 * D)
    out = torch.empty(M, 3 * D, dtype=odt, device="cuda")
    ops.gemm(A.cuda(), B.cuda(), bias.cuda(), out, 4, rope=freqs.cuda(), rope_cols=2 * D)
    tol = 1e-5 if odt == torch.float32 else 4e-3
    assert rel_l2(out.float(), want) < tol
    assert rel_l2(out.float()[:, 2 * D:], want[:, 2 * D:]) < tol


@pytest.mark.parametrize("wdt,odt", [(torch.float32, torch.float32), (torch.bfloat16, torch.bfloat16), (torch.bfloat16, torch.float32)])
@pytest.mark.parametrize("M,N,K", [(256, 1152, 3456), (200, 768, 2048), (1000, 3072, 1152), (512, 128, 64), (136, 2736, 1024)])
def test_gemm_dgrad_b_mn(ops, wdt, odt, M, N, K):
    """dX[M,N] = dY[M,K] W[K,N]: the weight is read in its nn.Linear layout through b_mn (no transposed copy)."""
    g = torch.Generator().manual_seed(M + N)
    dY = torch.randn(M, K, generator=g).to(wdt)
    W = (torch.randn(K, N, generator=g) / K ** 0.5).to(wdt)
    want = dY.double() @ W.double()
    out = torch.empty(M, N, dtype=odt, device="cuda")
    ops.gemm(dY.cuda(), W.cuda(), None, out, EPI_BIAS, b_mn=True)
    assert rel_l2(out.float(), want) < (1e-5 if odt == torch.float32 else 4e-3)


@pytest.mark.parametrize("wdt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rows,N,K", [(512, 1152, 1152), (400, 3456, 768), (8192, 256, 128), (77 * 8, 2736, 1024), (200, 16, 768)])
def test_gemm_wgrad_ab_mn(ops, wdt, rows, N, K):
    """dW[N,K] = dY[rows,N]^T X[rows,K]: both operands contraction-major (a_mn, b_mn), contraction over the token rows, fp32 out."""
    g = torch.Generator().manual_seed(rows + N)
    dY = torch.randn(rows, N, generator=g).to(wdt)
    X = torch.randn(rows, K, generator=g).to(wdt)
    want = dY.double().t() @ X.double()
    out = torch.empty(N, K, dtype=torch.float32, device="cuda")
    ops.gemm(dY.cuda(), X.cuda(), None, out, EPI_BIAS, a_mn=True, b_mn=True)
    assert rel_l2(out, want) < 1e-5


# ---------------------------------------------------------------------------------------------- LayerNorm + modulate prologue (fit.py:45-46,332,336)
def _ln_case(M, N, K, rps, seed, mean_shift=0.0):
    g = torch.Generator().manual_seed(seed)
    S = (M + rps - 1) // rps
    x = (torch.randn(M, K, generator=g) * (0.5 + torch.rand(M, 1, generator=g)) + mean_shift * torch.randn(M, 1, generator=g)).bfloat16()
    W = (torch.randn(N, K, generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(N, generator=g) * 0.1
    mod = torch.randn(S, 3 * K, generator=g) * 0.3     # shift | scale | (gate: unused) -- strided views like the adaLN table
    return x, W, bias, mod, S


def _ln_reference(x, W, bias, mod, rps, K):
    """modulate(LayerNorm(x), shift, scale) @ W^T + bias in fp64 on the same bf16 inputs."""
    xd = x.double()
    h = F.layer_norm(xd, (K,), eps=1e-6)
    S = mod.shape[0]
    sh = mod[:, :K].double().repeat_interleave(rps, 0)[:x.shape[0]]
    sc = mod[:, K:2 * K].double().repeat_interleave(rps, 0)[:x.shape[0]]
    return (h * (1 + sc) + sh) @ W.double().t() + bias.double()


def _ln_operands(ops, x, W, bias, mod, S, K, N):
    """What FiTEngine prepares: packed modulation rows, the (c1, c2) GEMM, the row sums."""
    modc = mod.cuda()
    packed = torch.empty(1, 2 * S, K, dtype=torch.bfloat16, device="cuda")
    ops.modulation_pack(modc, modc[:, K:], 3 * K, S, K, 1, 3 * K, packed)
    c = torch.empty(2 * S, N, dtype=torch.float32, device="cuda")
    ops.gemm(packed[0], W.cuda(), bias.cuda(), c, EPI_BIAS, bias_row0=S)
    parts = torch.full((x.shape[0], 3, 2), float("nan"), dtype=torch.float32, device="cuda")
    ops.row_stats(x.cuda(), parts)
    return packed[0, :S], c, parts


LN_SHAPES = [(512, 3456, 1152, 256),      # the qkv shape: one sample per 256-row pair tile
             (1024, 768, 768, 128),        # FiT-B width, one sample per CTA
             (400, 432, 192, 256),         # ragged last row block and last sample, 3 k-slabs
             (1000, 520, 1152, 128),       # cut last column tile, last row block partly beyond M
             (38000, 256, 128, 384)]       # several tiles per CTA pair: ring / TMEM phases wrap under the A-producer warps


@pytest.mark.parametrize("M,N,K,rps", LN_SHAPES)
def test_gemm_ln_prologue_bias(ops, M, N, K, rps):
    x, W, bias, mod, S = _ln_case(M, N, K, rps, M + N)
    want = _ln_reference(x, W, bias, mod, rps, K)
    g, c, parts = _ln_operands(ops, x, W, bias, mod, S, K, N)
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    ops.gemm(x.cuda(), W.cuda(), None, out, EPI_BIAS, rows_per_sample=rps, ln=(g, c, parts, 1e-6))
    # the modulated operand is rounded to bf16 twice over (x * g, then the fp32 epilogue): same order as the stand-alone kernel's single rounding
    assert rel_l2(out.float(), want) < 6e-3
    # against the un-fused kernels on the same inputs: same result up to operand rounding
    h = torch.empty(M, K, dtype=torch.bfloat16, device="cuda")
    modc = mod.cuda()
    ops.ln_modulate(x.cuda(), modc, modc[:, K:], 3 * K, rps, h)
    ref = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    ops.gemm(h, W.cuda(), bias.cuda(), ref, EPI_BIAS)
    assert rel_l2(out.float(), ref.float()) < 8e-3


def test_gemm_ln_prologue_row_mean(ops):
    """Rows whose mean is several standard deviations: the mean is removed in the epilogue (mean * c1), not on the operand."""
    M, N, K, rps = 512, 512, 1152, 256
    x, W, bias, mod, S = _ln_case(M, N, K, rps, 3, mean_shift=3.0)
    want = _ln_reference(x, W, bias, mod, rps, K)
    g, c, parts = _ln_operands(ops, x, W, bias, mod, S, K, N)
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    ops.gemm(x.cuda(), W.cuda(), None, out, EPI_BIAS, rows_per_sample=rps, ln=(g, c, parts, 1e-6))
    assert rel_l2(out.float(), want) < 2e-2


@pytest.mark.parametrize("M,T,heads,hd", [(512, 256, 16, 72), (640, 128, 4, 64)])
def test_gemm_ln_prologue_rope(ops, M, T, heads, hd):
    D = heads * hd
    N, K = 3 * D, D
    x, W, bias, mod, S = _ln_case(M, N, K, T, M + T)
    want = _ln_reference(x, W, bias, mod, T, K)
    gen = torch.Generator().manual_seed(9)
    ang = torch.rand(1, T, hd // 2, generator=gen) * 6.28
    freqs = torch.stack([ang.cos(), ang.sin()], -1).reshape(1, T, hd).float().contiguous()
    tok = torch.arange(M) % T
    w2 = want[:, :2 * D].reshape(M, 2 * heads, hd // 2, 2)
    cs = freqs[0, tok].double().reshape(M, 1, hd // 2, 2)
    rot = torch.stack([w2[..., 0] * cs[..., 0] - w2[..., 1] * cs[..., 1], w2[..., 1] * cs[..., 0] + w2[..., 0] * cs[..., 1]], -1)
    want = torch.cat([rot.reshape(M, 2 * D), want[:, 2 * D:]], 1)
    g, c, parts = _ln_operands(ops, x, W, bias, mod, S, K, N)
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    ops.gemm(x.cuda(), W.cuda(), None, out, 4, rope=freqs.cuda(), rope_cols=2 * D, rows_per_sample=T, ln=(g, c, parts, 1e-6))
    assert rel_l2(out.float(), want) < 6e-3


@pytest.mark.parametrize("M,K,H,rps", [(512, 1152, 3072, 256), (300, 768, 2048, 128), (256, 1024, 2730, 256)])
def test_gemm_ln_prologue_swiglu(ops, M, K, H, rps):
    x, _, _, mod, S = _ln_case(M, 8, K, rps, M + H)
    gen = torch.Generator().manual_seed(H)
    Wg = (torch.randn(H, K, generator=gen) / K ** 0.5).bfloat16()
    Wx = (torch.randn(H, K, generator=gen) / K ** 0.5).bfloat16()
    bg, bx = torch.randn(H, generator=gen) * 0.1, torch.randn(H, generator=gen) * 0.1
    want = F.silu(_ln_reference(x, Wg, bg, mod, rps, K)) * _ln_reference(x, Wx, bx, mod, rps, K)
    Hp = (H + 15) // 16 * 16
    wg, wx = torch.zeros(Hp, K, dtype=torch.bfloat16), torch.zeros(Hp, K, dtype=torch.bfloat16)
    wg[:H], wx[:H] = Wg, Wx
    b1, b2 = torch.zeros(Hp), torch.zeros(Hp)
    b1[:H], b2[:H] = bg, bx
    Wp = torch.stack([wg.view(Hp // 16, 16, K), wx.view(Hp // 16, 16, K)], 1).reshape(2 * Hp, K).contiguous()
    bp = torch.stack([b1.view(Hp // 16, 16), b2.view(Hp // 16, 16)], 1).reshape(2 * Hp).contiguous()
    g, c, parts = _ln_operands(ops, x, Wp, bp, mod, S, K, 2 * Hp)
    out = torch.empty(M, Hp, dtype=torch.bfloat16, device="cuda")
    ops.gemm(x.cuda(), Wp.cuda(), None, out, EPI_SWIGLU, rows_per_sample=rps, ln=(g, c, parts, 1e-6))
    assert rel_l2(out[:, :H].float(), want) < 8e-3
    assert float(out[:, H:].float().abs().sum()) == 0.0


@pytest.mark.parametrize("adt", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("M,N,K,rps", [(512, 1152, 1152, 256), (13200, 864, 192, 200), (300, 128, 64, 100)])
def test_gemm_gate_residual_row_sums(ops, adt, M, N, K, rps):
    """stats_out of the GATE_RESID epilogue: the partial sums add up to sum / sum of squares of every output row."""
    A, B, bias = _inputs(M, N, K, M + N + 1, torch.bfloat16)
    gen = torch.Generator().manual_seed(2)
    R = (M + rps - 1) // rps
    gate = torch.randn(R, N, generator=gen).cuda()
    xs = torch.randn(M, N, generator=gen).to(adt).cuda()
    parts = torch.full((M, ops.gemm_stats_parts(N), 2), float("nan"), dtype=torch.float32, device="cuda")
    ops.gemm(A.cuda(), B.cuda(), bias.cuda(), xs, EPI_BIAS_GATE_RESID, gate=gate, ld_gate=N, rows_per_sample=rps, resid=xs, stats_out=parts)
    s = parts.double().sum(1)
    xd = xs.double()
    tol = 2e-2 if adt == torch.bfloat16 else 1e-4   # the sums are of the fp32 values before the bf16 store
    assert float((s[:, 0] - xd.sum(1)).abs().max()) < tol * float(xd.abs().sum(1).max())
    assert rel_l2(s[:, 1].float(), (xd * xd).sum(1)) < tol


def test_gemm_ln_prologue_argument_errors(ops):
    x, W, bias, mod, S = _ln_case(400, 256, 128, 200, 1)
    g, c, parts = _ln_operands(ops, x, W, bias, mod, S, 128, 256)
    out = torch.empty(400, 256, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(RuntimeError, match="rows_per_sample"):   # CTA tiles would straddle samples: callers take the ln_modulate kernel
        ops.gemm(x.cuda(), W.cuda(), None, out, EPI_BIAS, rows_per_sample=200, ln=(g, c, parts, 1e-6))


@pytest.mark.parametrize("Nout,Kin,tokens", [(1152, 1152, 8192), (768, 768, 4096), (1152, 3072, 8192)])
def test_wgrad_split_k_matches_unsplit(ops, Nout, Kin, tokens, monkeypatch):
    """Weight-gradient GEMMs with few output tiles split the token contraction into work items whose partial tiles are added into the zeroed fp32
    output by TMA reduce stores (cost model in launch_tc; dW_proj at 8192 rows: 2 splits).  Against fp64 math on the same inputs, with a
    NaN-filled output buffer (the entry point zeroes it itself)."""
    g = torch.Generator().manual_seed(Nout + Kin)
    dy = torch.randn(tokens, Nout, generator=g).bfloat16()
    x = torch.randn(tokens, Kin, generator=g).bfloat16()
    want = dy.double().t() @ x.double()
    out = torch.full((Nout, Kin), float("nan"), device="cuda")
    ops.gemm(dy.cuda(), x.cuda(), None, out, EPI_BIAS, a_mn=True, b_mn=True)
    assert rel_l2(out, want) < 1e-5
    # a strided output (a view into a gradient bucket row): only the view's columns are zeroed / written
    big = torch.full((Nout, Kin + 64), 7.0, device="cuda")
    ops.gemm(dy.cuda(), x.cuda(), None, big[:, :Kin], EPI_BIAS, a_mn=True, b_mn=True)
    assert rel_l2(big[:, :Kin], want) < 1e-5 and float((big[:, Kin:] - 7.0).abs().max()) == 0.0
