"""GPU parity of the BASELINE.json configurations the first round left untested ("Next round" item 1):
  * config 4 at FiT-XL/2 width: loss + sampled parameter gradients vs the reference's goldens (tests/golden/fit_xl2_small.npz);
  * config 5, masked leg: masked_FiT batch assembly on device (bit-exact vs the reference's own collate) + the training step on the
    sub-sampled tokens with their permuted RoPE rows vs the reference's loss / gradients (tests/golden/masked_fit_tiny.npz);
  * config 5, extrapolation leg at 1024x1024: 4096-token attention (attn_tc3: 32 key blocks per row, O rescaled in TMEM) vs the oracle.
The measured worst cases are recorded in gpurun_out/parity_measured.txt."""
import numpy as np
import pytest
import torch

from helpers import build_model, golden, record, rel_l2
from oracle import diffusion_oracle as do
from oracle import fit_oracle as fo
from oracle import synth

pytestmark = pytest.mark.gpu


def _grad_errors(z, model, sampled):
    errs = {}
    for name, p in model.named_parameters():
        assert p.grad is not None, name
        got = p.grad.float().cpu()
        ref = torch.tensor(z["g." + name])
        if sampled:
            f = got.reshape(-1)
            got = f[:: max(1, f.numel() // 2048)][:2048]
        # sampled entries are compared on the scale of the sample (floored by 0.1 % of the full-gradient norm for near-empty samples)
        den = max(float(ref.double().norm()), 1e-3 * float(z["gnorm." + name]), 1e-30)
        errs[name] = float((got.double().reshape(-1) - ref.double().reshape(-1)).norm()) / den
        if not sampled:
            gn = float(z["gnorm." + name])
            assert abs(float(p.grad.double().norm()) - gn) <= 0.1 * gn + 1e-12, name
    return errs


def _by_class(errs):
    """Worst error per tensor class: weights of the big GEMMs, biases, adaLN tables, embedders."""
    cls = {}
    for k, v in errs.items():
        c = ("adaLN" if "adaLN" in k else "embedder" if k.split(".")[0] in ("x_embedder", "t_embedder", "y_embedder") else
             "final" if k.startswith("final_layer") else "bias" if k.endswith("bias") else "weight")
        if v > cls.get(c, ("", 0.0))[1]:
            cls[c] = (k, v)
    return cls


# bf16 bounds per tensor class (measured worst cases are recorded; a mis-scaled reduction in any tensor shows up as an error of order 1)
BF16_GRAD_TOL = {"weight": 4e-2, "bias": 4e-2, "adaLN": 5e-2, "embedder": 5e-2, "final": 4e-2}


@pytest.mark.parametrize("precision,tol_loss,tol_grad", [("fp32", 1e-5, 2e-3), ("bf16", 2e-2, None)])
def test_xl2_width_training_step_matches_reference(precision, tol_loss, tol_grad):
    """BASELINE config 4 shape per row: FiT-XL/2 (D = 1152, H = 3072, 16 heads of 72, 28 blocks), two mixed-resolution rows padded to
    256 tokens: loss and 2048 sampled entries of every parameter gradient vs the unmodified reference (attn_bwd_tc<72>, the
    contraction-major pair-tile dgrad / wgrad GEMMs at K = 1152 / 3072)."""
    from fit_b200.train import masked_mse_loss
    z = golden("fit_xl2_small")
    cfg = fo.model_config("FiT-XL/2")
    sd = fo.random_state_dict(cfg, seed=7)
    x, t, y, pos, mask, target = synth.mixed_batch(cfg, [(32, 32), (20, 40)], 256, seed=11, rope_dim=72)
    model = build_model(cfg, sd, precision, train_mode=True)
    model.y_embedder.dropout_prob = 0.0
    out = model(x.cuda(), t.cuda(), y.cuda(), pos.cuda(), mask.cuda(), train=True)
    loss = masked_mse_loss(out, target.cuda(), mask.cuda())
    loss.backward()
    assert abs(float(loss) - float(z["loss"])) < tol_loss * abs(float(z["loss"]))
    errs = _grad_errors(z, model, sampled=True)
    cls = _by_class(errs)
    record(f"xl2_train_step[{precision}] loss_rel", abs(float(loss) - float(z["loss"])) / abs(float(z["loss"])))
    for c, (k, v) in cls.items():
        record(f"xl2_train_step[{precision}] worst grad rel-L2 {c}", f"{v:.3e} ({k})")
        assert v < (tol_grad if tol_grad is not None else BF16_GRAD_TOL[c]), (c, k, v)


@pytest.mark.parametrize("precision,tol_loss,tol_grad", [("fp32", 1e-5, 1e-3), ("bf16", 2e-2, None)])
def test_masked_fit_training_step_matches_reference(precision, tol_loss, tol_grad):
    """masked_FiT (config 5, masked leg; row f4): token sub-sampling on the device == the reference's own collate bit for bit, then
    q_sample + forward + masked MSE + backward on the T' = 128 kept tokens (permuted order, per-sample RoPE rows) vs the reference."""
    from fit_b200.diffusion import create_diffusion
    from fit_b200.masked import subsample_batch
    from fit_b200.train import add_noise, masked_mse_loss
    z = golden("masked_fit_tiny")
    toks, poss, masks, labels = synth.masked_items()
    T2 = int(z["number_of_tokens"])
    lat, pos, msk = subsample_batch(toks, poss, masks, T2, "cuda", generator=torch.Generator().manual_seed(11))
    assert torch.equal(lat.cpu(), torch.from_numpy(z["collate.latent"]))
    assert torch.equal(pos.cpu(), torch.from_numpy(z["collate.pos"]))
    assert torch.equal(msk.cpu(), torch.from_numpy(z["collate.mask"]))
    t, noise = synth.masked_step_noise(lat.shape)
    x_t = add_noise(create_diffusion(""), lat, noise.cuda(), t.cuda())
    assert rel_l2(x_t, z["x_t"]) < 1e-6
    cfg = synth.MASKED_CFG
    sd = fo.random_state_dict(cfg, seed=7)
    model = build_model(cfg, sd, precision, train_mode=True)
    model.y_embedder.dropout_prob = 0.0
    out = model(x_t, t.cuda(), torch.tensor(labels).cuda(), pos, msk, train=True)
    loss = masked_mse_loss(out, noise.cuda(), msk)
    loss.backward()
    m = msk.cpu()
    assert rel_l2(out.detach().cpu()[m], torch.from_numpy(z["out_train"])[m]) < (1e-4 if precision == "fp32" else 2e-2)
    assert abs(float(loss) - float(z["loss"])) < tol_loss * abs(float(z["loss"]))
    errs = _grad_errors(z, model, sampled=False)
    for c, (k, v) in _by_class(errs).items():
        record(f"masked_train_step[{precision}] worst grad rel-L2 {c}", f"{v:.3e} ({k})")
        assert v < (tol_grad if tol_grad is not None else BF16_GRAD_TOL[c]), (c, k, v)


@pytest.mark.parametrize("ramp", [False, True])
def test_attention_4096_tokens(ramp):
    """attn_tc3 at the 1024 x 1024 extrapolation length: (T, h, d) = (4096, 16, 72), key counts around the last block boundary
    (4096, 4095, 3969 = 31 x 128 + 1, 2049).  `ramp`: keys grow along the sequence so the running row maximum rises block after block and
    the lazy O rescale (threshold 2^8) fires many times per row."""
    from fit_b200 import ops
    from test_gpu_kernels import _attn_case
    T, h, d = 4096, 16, 72
    lens = [4096, 4095, 3969, 2049]
    R = len(lens)
    qkv, freqs, mask = _attn_case(R, T, h, d, lens, seed=41)
    if ramp:
        qkv[:, :, 1] *= torch.linspace(0.25, 6.0, T)[None, :, None, None]
        qkv[:, :, 0] *= 2.0
    q3, k3, v3 = qkv.bfloat16().float().permute(2, 0, 3, 1, 4)
    qr, kr = fo.rope_rotate(q3, freqs).bfloat16().float(), fo.rope_rotate(k3, freqs).bfloat16().float()
    qkv_rot = torch.stack([qr, kr, v3], 0).permute(1, 3, 0, 2, 4).contiguous().bfloat16()
    kv, _ = ops.mask_to_kvlen(mask.cuda())
    out = torch.full((R * T, h * d), float("nan"), dtype=torch.bfloat16, device="cuda")
    lse = torch.full((R, h, T), float("nan"), device="cuda")
    ops.attention(qkv_rot.cuda().view(R * T, 3 * h * d), None, kv, mask.cuda(), R, T, h, d, out, lse, prerotated=True)
    assert bool(torch.isfinite(out.float()).all()) and bool(torch.isfinite(lse).all())
    out = out.float().cpu().view(R, T, h * d)
    worst, worst_lse = 0.0, 0.0
    for r, L in enumerate(lens):   # one sample at a time: the oracle's score tensor is h x T x T fp32 = 1 GB
        want = fo.attention_core(qr[r:r + 1], kr[r:r + 1], v3[r:r + 1], mask[r:r + 1]).transpose(1, 2).reshape(T, -1)
        sc = (qr[r] @ kr[r].transpose(-1, -2)) * d ** -0.5
        want_lse = torch.logsumexp(sc.masked_fill(~mask[r][None, None, :], float("-inf")), -1)
        del sc
        e = rel_l2(out[r, :L], want[:L])
        el = float((lse[r, :, :L].cpu() - want_lse[:, :L]).abs().max())
        worst, worst_lse = max(worst, e), max(worst_lse, el)
        assert e < 5e-3, (r, L, e)
        assert el < 3e-2, (r, L, el)
        q0 = (L + 127) // 128 * 128
        if q0 < T:
            assert float(out[r, q0:].abs().max()) == 0.0
    record(f"attention T=4096 h=16 d=72 ramp={ramp} worst rel-L2 / lse abs", f"{worst:.3e} / {worst_lse:.3e}")
