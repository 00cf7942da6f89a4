"""GPU parity of the training step: every backward kernel against torch autograd of the oracle's
formula on the same inputs (fp64 checker), then the whole network's loss and parameter gradients against the goldens
generated from the unmodified reference (tests/golden/fit_tiny_*.npz, fit_b2_mixed.npz).
Tolerances: fp32 verification mode 1e-3 rel-L2 per gradient (atomics / summation order), bf16 5e-2."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from helpers import TINY64, TINY72, TINY_SHAPES, build_model, golden, rel_l2

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
    assert rel_l2(db, dy.double().cpu().sum(0)) < 1e-4


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
    assert rel_l2(db.cpu().view(Hp // 16, 2, 16)[:, 0].reshape(Hp), du[:, :, 0].reshape(M, Hp).double().sum(0)) < 1e-4


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
