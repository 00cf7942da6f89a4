"""GPU parity of the training step: every backward kernel against torch autograd of the oracle's
formula on the same inputs (fp64 checker), then the whole network's loss and parameter gradients against the goldens
generated from the unmodified reference (tests/golden/fit_tiny_*.npz, fit_b2_mixed.npz).
Tolerances: fp32 verification mode 1e-3 rel-L2 per gradient (atomics / summation order), bf16 5e-2."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from helpers import TINY64, TINY72, TINY_SHAPES, build_model, golden, record, rel_l2

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from fit_b200 import ops as _ops
    return _ops


def _tol(dt, f32=1e-5, bf16=8e-3):
    return f32 if dt == torch.float32 else bf16


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("R,T,D", [(3, 64, 128), (2, 200, 768), (4, 33, 1152)])
def test_gate_resid_and_gate_bwd(ops, dt, R, T, D):
    g = torch.Generator().manual_seed(R * T)
    M = R * T
    resid, y, dx = (torch.randn(M, D, generator=g).to(dt) for _ in range(3))
    mod = torch.randn(R, 3 * D, generator=g)
    gate = mod[:, D:2 * D]
    gr = gate.double().repeat_interleave(T, 0)
    modc = mod.cuda()
    out = ops.gate_resid_fwd(resid.cuda(), y.cuda(), modc[:, D:2 * D], 3 * D, T, torch.empty(M, D, dtype=dt, device="cuda"))
    assert rel_l2(out.float(), resid.double() + gr * y.double()) < _tol(dt)
    dmod = torch.zeros(R, 3 * D, device="cuda")
    db = torch.zeros(D, device="cuda")
    dy = ops.gate_bwd(dx.cuda(), y.cuda(), modc[:, D:2 * D], 3 * D, R, T, torch.empty(M, D, dtype=dt, device="cuda"), dmod[:, D:2 * D], 3 * D, db)
    want_dy = gr * dx.double()
    assert rel_l2(dy.float(), want_dy) < _tol(dt)
    assert rel_l2(dmod[:, D:2 * D], (dx.double() * y.double()).view(R, T, D).sum(1)) < 1e-5
    assert float(dmod[:, :D].abs().sum()) == 0.0 and float(dmod[:, 2 * D:].abs().sum()) == 0.0
    # the bias gradient sums the values the kernel wrote (rounded to dt)
    assert rel_l2(db, want_dy.sum(0)) < 1e-4   # accumulated in fp32 before the rounding to dt


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("M,H", [(100, 2048), (64, 2730), (37, 336)])
def test_swiglu_fwd_bwd(ops, dt, M, H):
    Hp = (H + 15) // 16 * 16
    g = torch.Generator().manual_seed(M + H)
    ug, ux, dg = (torch.randn(M, Hp, generator=g).to(dt) for _ in range(3))
    u = torch.stack([ug.view(M, Hp // 16, 16), ux.view(M, Hp // 16, 16)], 2).reshape(M, 2 * Hp).contiguous()
    a = ug.double().requires_grad_(True)
    b = ux.double().requires_grad_(True)
    want = F.silu(a) * b
    want.backward(dg.double())
    got = ops.swiglu_fwd(u.cuda(), torch.empty(M, Hp, dtype=dt, device="cuda"))
    assert rel_l2(got.float(), want.detach()) < _tol(dt)
    db = torch.zeros(2 * Hp, device="cuda")
    du = ops.swiglu_bwd(dg.cuda(), u.cuda(), torch.empty(M, 2 * Hp, dtype=dt, device="cuda"), db).float().cpu().view(M, Hp // 16, 2, 16)
    assert rel_l2(du[:, :, 0].reshape(M, Hp), a.grad) < _tol(dt)
    assert rel_l2(du[:, :, 1].reshape(M, Hp), b.grad) < _tol(dt)
    assert rel_l2(db.cpu().view(Hp // 16, 2, 16)[:, 0].reshape(Hp), a.grad.sum(0)) < 1e-4   # fp32 sums of the unrounded values
    assert rel_l2(db.cpu().view(Hp // 16, 2, 16)[:, 1].reshape(Hp), b.grad.sum(0)) < 1e-4


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("R,T,D", [(2, 256, 1152), (3, 200, 768), (4, 17, 144)])
def test_ln_modulate_bwd(ops, dt, R, T, D):
    from oracle import fit_oracle as fo
    g = torch.Generator().manual_seed(D + T)
    M = R * T
    x = (torch.randn(M, D, generator=g) * 2 + 0.3).to(dt)
    dh, dres = torch.randn(M, D, generator=g).to(dt), torch.randn(M, D, generator=g).to(dt)
    mod = torch.randn(R, 2 * D, generator=g) * 0.5
    xd = x.double().requires_grad_(True)
    sh = mod[:, :D].double().requires_grad_(True)
    sc = mod[:, D:].double().requires_grad_(True)
    hh = fo.modulate(F.layer_norm(xd.view(R, T, D), (D,), eps=1e-6), sh, sc)
    hh.backward(dh.double().view(R, T, D))
    modc = mod.cuda()
    mean, rstd = torch.empty(M, device="cuda"), torch.empty(M, device="cuda")
    h = ops.ln_modulate(x.cuda(), modc[:, :D], modc[:, D:], 2 * D, T, torch.empty(M, D, dtype=dt, device="cuda"), mean, rstd)
    assert rel_l2(h.float(), hh.detach().view(M, D)) < _tol(dt)
    dmod = torch.zeros(R, 2 * D, device="cuda")
    dx = ops.ln_modulate_bwd(dh.cuda(), x.cuda(), mean, rstd, modc[:, D:], 2 * D, T, dres.cuda(), torch.empty(M, D, dtype=dt, device="cuda"),
                             dmod[:, :D], dmod[:, D:], 2 * D)
    assert rel_l2(dx.float(), xd.grad + dres.double()) < _tol(dt, 2e-5)
    assert rel_l2(dmod[:, :D], sh.grad) < 1e-4 and rel_l2(dmod[:, D:], sc.grad) < 1e-4
    dx2 = ops.ln_modulate_bwd(dh.cuda(), x.cuda(), mean, rstd, modc[:, D:], 2 * D, T, None, torch.empty(M, D, dtype=dt, device="cuda"),
                              None, None, 2 * D)
    assert rel_l2(dx2.float(), xd.grad) < _tol(dt, 2e-5)


def test_small_backward_pieces(ops):
    g = torch.Generator().manual_seed(5)
    a = torch.randn(300, 200, generator=g)
    out = ops.colsum(a.cuda()[:, :136], torch.zeros(136, device="cuda"))      # strided view
    assert rel_l2(out, a[:, :136].double().sum(0)) < 1e-5
    outb = ops.colsum(a.bfloat16().cuda(), torch.zeros(200, device="cuda"))
    assert rel_l2(outb, a.bfloat16().double().sum(0)) < 1e-5
    # vectorised path at a training shape (ragged last row chunk), the thread-per-column path for widths / strides that are not multiples of 8,
    # accumulation into a non-zero output
    big = torch.randn(8200, 3456, generator=g).bfloat16()
    outv = ops.colsum(big.cuda(), torch.ones(3456, device="cuda"))
    assert rel_l2(outv, big.double().sum(0) + 1.0) < 1e-5
    odd = torch.randn(300, 203, generator=g)
    outo = ops.colsum(odd.cuda()[:, :13], torch.zeros(13, device="cuda"))
    assert rel_l2(outo, odd[:, :13].double().sum(0)) < 1e-5
    x = torch.randn(7, 144, generator=g).double().requires_grad_(True)
    dy = torch.randn(7, 144, generator=g)
    F.silu(x).backward(dy.double())
    y = ops.silu_fwd(x.detach().float().cuda(), torch.empty(7, 144, device="cuda"))
    assert rel_l2(y, F.silu(x.detach())) < 1e-6
    dx = ops.silu_bwd(dy.cuda(), x.detach().float().cuda(), torch.empty(7, 144, device="cuda"))
    assert rel_l2(dx, x.grad) < 1e-6
    lab = torch.tensor([3, 0, 3, 10, 7, 3, 0])
    dt_ = ops.embedding_bwd(dy.cuda(), lab.cuda(), torch.zeros(11, 144, device="cuda"))
    want = torch.zeros(11, 144, dtype=torch.float64).index_add_(0, lab, dy.double())
    assert rel_l2(dt_, want) < 1e-6
    b = ops.cast(a.cuda(), torch.empty(300, 200, dtype=torch.bfloat16, device="cuda"))
    assert torch.equal(b.cpu(), a.bfloat16())


def test_masked_mse(ops):
    from fit_b200.train import masked_mse_loss
    from oracle import fit_oracle as fo
    g = torch.Generator().manual_seed(9)
    out = torch.randn(5, 64, 16, generator=g)
    tgt = torch.randn(5, 64, 16, generator=g)
    mask = torch.zeros(5, 64, dtype=torch.bool)
    for r, L in enumerate([64, 10, 33, 1, 50]):
        mask[r, :L] = True
    o = out.double().requires_grad_(True)
    want = fo.masked_mse(o, tgt.double(), mask) * 3.0
    want.backward()
    oc = out.cuda().requires_grad_(True)
    loss = masked_mse_loss(oc, tgt.cuda(), mask.cuda()) * 3.0
    loss.backward()
    assert abs(float(loss) - float(want)) < 1e-5 * abs(float(want))
    assert rel_l2(oc.grad, o.grad) < 1e-6


# ------------------------------------------------------------------------------ attention backward
@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("T,h,d,lens,nonprefix", [
    (64, 2, 64, [64, 50, 36], True),
    (96, 3, 72, [96, 10, 33], False),
    (256, 2, 72, [256, 200, 129, 17], False),
    (200, 2, 64, [200, 7], True),
    (256, 3, 64, [256, 128, 130, 16], False),
    (8, 1, 72, [8, 3], False),
])
def test_attention_bwd(ops, dt, T, h, d, lens, nonprefix):
    from oracle import fit_oracle as fo
    from test_gpu_kernels import _attn_case
    R = len(lens)
    qkv, freqs, mask = _attn_case(R, T, h, d, lens, seed=T + d + 1, nonprefix=nonprefix)
    g = torch.Generator().manual_seed(3)
    dO = torch.randn(R, T, h * d, generator=g)
    for r, L in enumerate(lens):   # the loss is masked: no gradient flows into padded query rows (FiT/train.py:63-66)
        dO[r, L:] = 0
    if dt == torch.bfloat16:
        qkv, dO = qkv.bfloat16().float(), dO.bfloat16().float()
    leaf = qkv.double().requires_grad_(True)
    q, k, v = leaf.permute(2, 0, 3, 1, 4)
    fr = freqs.double()
    o = fo.attention_core(fo.rope_rotate(q, fr), fo.rope_rotate(k, fr), v, mask).transpose(1, 2).reshape(R, T, h * d)
    o.backward(dO.double())
    want = leaf.grad.reshape(R, T, 3 * h * d)
    kv, _ = ops.mask_to_kvlen(mask.cuda())
    tol = 2e-5 if dt == torch.float32 else 1.5e-2
    # (pre-rotated operands, CUDA-core kernel forced): bf16 + pre-rotated defaults to the tcgen05 kernel for T <= 256
    # bf16 forward operands always carry the RoPE (fitb200_attn_fwd refuses un-rotated bf16)
    variants = [(False, False)] if dt == torch.float32 else [(True, False), (True, True)]
    for pre, simt in variants:
        if pre:   # forward operands pre-rotated (the bf16 path); the gradient is still for the un-rotated q, k
            q3, k3, v3 = qkv.permute(2, 0, 3, 1, 4)
            src = torch.stack([fo.rope_rotate(q3, freqs), fo.rope_rotate(k3, freqs), v3], 0).permute(1, 3, 0, 2, 4).contiguous()
        else:
            src = qkv
        src_c = src.to(dt).cuda().view(R * T, 3 * h * d)
        out = torch.empty(R * T, h * d, dtype=dt, device="cuda")
        lse = torch.empty(R, h, T, device="cuda")
        ops.attention(src_c, None if pre else freqs.cuda(), kv, mask.cuda(), R, T, h, d, out, lse, prerotated=pre)
        dqkv = torch.full((R * T, 3 * h * d), float("nan"), dtype=dt, device="cuda")
        ops.attention_bwd(src_c, freqs.cuda(), kv, mask.cuda(), R, T, h, d, out, dO.to(dt).cuda().view(R * T, h * d), lse, dqkv, prerotated=pre,
                          force_simt=simt)
        got = dqkv.float().cpu().view(R, T, 3 * h * d)
        assert bool(torch.isfinite(got).all())
        for r, L in enumerate(lens):
            assert rel_l2(got[r, :L], want[r, :L]) < tol, (pre, simt, r, L)
            assert float(got[r, L:].abs().max()) == 0.0 if L < T and not nonprefix else True


# ------------------------------------------------------------------------------ whole network: loss + every parameter gradient
def _grad_check(z, model, sampled, tol):
    worst = ("", 0.0)
    for name, p in model.named_parameters():
        assert p.grad is not None, name
        got = p.grad.float().cpu()
        ref = torch.tensor(z["g." + name])
        if sampled or tuple(ref.shape) != tuple(got.shape):
            f = got.reshape(-1)
            got = f[:: max(1, f.numel() // 2048)][:2048]
            # sampled entries: compare against the full-gradient norm scale
            err = float((got.double() - ref.double()).norm() / max(float(ref.double().norm()), 1e-30))
        else:
            err = rel_l2(got, ref)
            assert abs(float(got.double().norm()) - float(z["gnorm." + name])) <= tol * float(z["gnorm." + name]) + 1e-12, name
        if err > worst[1]:
            worst = (name, err)
    return worst


@pytest.mark.parametrize("name,cfg", [("fit_tiny_d64", TINY64), ("fit_tiny_d72", TINY72)])
# measured on B200 (profiles/*_parity_measured.txt): fp32 worst gradient 7.5e-7, bf16 8.1e-3 (loss 4e-5)
@pytest.mark.parametrize("precision,tol_loss,tol_grad", [("fp32", 1e-5, 2e-5), ("bf16", 1e-3, 2e-2)])
def test_tiny_training_step_matches_reference(name, cfg, precision, tol_loss, tol_grad):
    """loss.backward() through the custom autograd node vs the reference's autograd (goldens from the unmodified reference)."""
    from fit_b200.train import masked_mse_loss
    from oracle import fit_oracle as fo
    from oracle import synth
    z = golden(name)
    sd = fo.random_state_dict(cfg, seed=7)
    d = cfg["hidden_size"] // cfg["num_heads"]
    x, t, y, pos, mask, target = synth.mixed_batch(cfg, TINY_SHAPES, 64, seed=11, rope_dim=d)
    model = build_model(cfg, sd, precision, train_mode=True)
    model.y_embedder.dropout_prob = 0.0   # label dropout off: the golden was generated without it (eval-mode embedder)
    out = model(x.cuda(), t.cuda(), y.cuda(), pos.cuda(), mask.cuda(), train=True)
    assert out.requires_grad
    loss = masked_mse_loss(out, target.cuda(), mask.cuda())
    loss.backward()
    assert abs(float(loss) - float(z["loss"])) < tol_loss * abs(float(z["loss"]))
    worst = _grad_check(z, model, sampled=False, tol=tol_grad)
    record(f"tiny_train_step {name} [{precision}] loss_rel / worst grad rel-L2", f"{abs(float(loss) - float(z['loss'])) / abs(float(z['loss'])):.3e} / {worst[1]:.3e} ({worst[0]})")
    assert worst[1] < tol_grad, worst


# measured on B200: fp32 worst gradient 6.5e-6, bf16 1.5e-2 (12 blocks; loss 2e-5)
@pytest.mark.parametrize("precision,tol_loss,tol_grad", [("fp32", 1e-5, 1e-4), ("bf16", 1e-3, 3e-2)])
def test_b2_mixed_resolution_training_step(precision, tol_loss, tol_grad):
    """BASELINE config 2: FiT-B/2, mixed-resolution batch (256^2, 160x320, 128x512 padded to 256 tokens + mask), forward + backward."""
    from fit_b200.train import masked_mse_loss
    from oracle import fit_oracle as fo
    from oracle import synth
    z = golden("fit_b2_mixed")
    cfg = fo.model_config("FiT-B/2")
    sd = fo.random_state_dict(cfg, seed=7)
    x, t, y, pos, mask, target = synth.mixed_batch(cfg, [(32, 32), (20, 40), (16, 64)], 256, seed=11, rope_dim=64)
    model = build_model(cfg, sd, precision, train_mode=True)
    model.y_embedder.dropout_prob = 0.0
    out = model(x.cuda(), t.cuda(), y.cuda(), pos.cuda(), mask.cuda(), train=True)
    loss = masked_mse_loss(out, target.cuda(), mask.cuda())
    loss.backward()
    assert abs(float(loss) - float(z["loss"])) < tol_loss * abs(float(z["loss"]))
    full = z["g.x_embedder.bias"].size == cfg["hidden_size"] and z["g.blocks.0.attn.qkv.weight"].size == 3 * cfg["hidden_size"] ** 2
    worst = _grad_check(z, model, sampled=not full, tol=tol_grad)
    record(f"b2_mixed_train_step [{precision}] loss_rel / worst grad rel-L2", f"{abs(float(loss) - float(z['loss'])) / abs(float(z['loss'])):.3e} / {worst[1]:.3e} ({worst[0]})")
    assert worst[1] < tol_grad, worst
