"""Pin the CPU oracle against fixtures produced by the unmodified reference
(oracle/make_goldens.py).  CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import diffusion_oracle as do
from oracle import fit_oracle as fo
from oracle import synth
from oracle.pos_embed_oracle import precompute_freqs_cis_2d

G = os.path.join(os.path.dirname(__file__), "golden")


def rel_l2(a, b):
    a, b = torch.as_tensor(a).double(), torch.as_tensor(b).double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def checksum(t):
    return float(t.double().abs().sum())


def sample_flat(t, n=2048):
    f = t.reshape(-1)
    step = max(1, f.numel() // n)
    return f[::step][:n]


def test_rope_tables_match_reference():
    z = np.load(os.path.join(G, "rope_tables.npz"))
    for key in z.files:
        parts = key.split("_")
        dim, nh, nw = int(parts[0][1:]), int(parts[1][1:]), int(parts[2][1:])
        ml = None if parts[3] == "mlNone" else int(parts[3][2:])
        tab = precompute_freqs_cis_2d(dim, nh, nw, max_length=ml)
        if key.endswith("_sample"):
            got = tab.reshape(-1)[:: tab.size // 4096][:4096]
            np.testing.assert_array_equal(got, z[key])
        elif key.endswith("_sum"):
            assert np.abs(tab).sum() == z[key]
        else:
            assert tab.shape == (nh * nw, dim)
            np.testing.assert_array_equal(tab, z[key])   # bit-exact: same numpy ops


def test_rope_layout_w_first_then_h():
    # first d/2 entries use the token's w coordinate, last d/2 its h coordinate (pos_embed.py:111-113)
    tab = precompute_freqs_cis_2d(64, 3, 5)
    tok = 1 * 5 + 4   # h=1, w=4
    f0 = 1.0
    assert abs(tab[tok, 0] - np.cos(4 * f0)) < 1e-6 and abs(tab[tok, 1] - np.sin(4 * f0)) < 1e-6
    assert abs(tab[tok, 32] - np.cos(1 * f0)) < 1e-6 and abs(tab[tok, 33] - np.sin(1 * f0)) < 1e-6


@pytest.mark.parametrize("rs", ["10", "250", "ddim25", ""])
def test_diffusion_tables_match_reference(rs):
    z = np.load(os.path.join(G, "diffusion_tables.npz"))
    tb, tmap = do.create(rs)
    tag = rs or "full"
    np.testing.assert_array_equal(np.array(tmap), z[tag + ".timestep_map"])
    for f in ["betas", "alphas_cumprod", "alphas_cumprod_prev", "sqrt_alphas_cumprod",
              "sqrt_one_minus_alphas_cumprod", "sqrt_recip_alphas_cumprod", "sqrt_recipm1_alphas_cumprod",
              "posterior_variance", "posterior_log_variance_clipped", "posterior_mean_coef1",
              "posterior_mean_coef2"]:
        np.testing.assert_array_equal(getattr(tb, f), z[tag + "." + f], err_msg=f)


def test_diffusion_known_values():
    tb, tmap = do.create("250")
    assert tmap[:3] == [0, 4, 8] and tmap[-2:] == [995, 999] and len(tmap) == 250
    assert abs(tb.sqrt_recip_alphas_cumprod[-1] - 157.41) < 0.01
    tb10, tmap10 = do.create("10")
    assert tmap10 == [0, 111, 222, 333, 444, 555, 666, 777, 888, 999]


def test_diffusion_steps_match_reference():
    z = np.load(os.path.join(G, "diffusion_steps.npz"))
    tb, _ = do.create("10")
    x, eps, noise = (torch.from_numpy(z[k]) for k in ("x", "eps", "noise"))
    t = torch.from_numpy(z["t"])
    o = do.p_sample(tb, x, t, eps, noise)
    np.testing.assert_array_equal(o["sample"].numpy(), z["p_sample.sample"])
    np.testing.assert_array_equal(o["pred_xstart"].numpy(), z["p_sample.pred_xstart"])
    pm = do.p_mean_variance(tb, x, t, eps)
    for k in ("mean", "variance", "log_variance"):
        np.testing.assert_array_equal(pm[k].numpy(), z["pmv." + k])
    for eta in (0.0, 0.5):
        o = do.ddim_sample(tb, x, t, eps, noise, eta=eta)
        np.testing.assert_array_equal(o["sample"].numpy(), z[f"ddim{eta}.sample"])
    np.testing.assert_array_equal(do.q_sample(tb, x, t, noise).numpy(), z["q_sample"])


TINY64 = dict(patch_size=2, in_channels=4, hidden_size=128, depth=2, num_heads=2, mlp_ratio=4.0,
              class_dropout_prob=0.1, num_classes=10, learn_sigma=False)
TINY72 = dict(TINY64, hidden_size=144)
TINY_SHAPES = [(16, 16), (10, 20), (8, 32), (12, 12)]


def _check_fit(name, cfg, shapes, T, tol, with_streams, grad_tol):
    z = np.load(os.path.join(G, name + ".npz"))
    sd = fo.random_state_dict(cfg, seed=7)
    d = cfg["hidden_size"] // cfg["num_heads"]
    x, t, y, pos, mask, target = synth.mixed_batch(cfg, shapes, T, seed=11, rope_dim=d)
    ck = np.array([checksum(x), checksum(pos), float(t.sum()), float(y.sum()), checksum(target),
                   sum(checksum(v) for v in sd.values())])
    np.testing.assert_allclose(ck, z["in_checksum"], rtol=1e-12)   # identical inputs and weights
    streams = [] if with_streams else None
    with torch.no_grad():
        out = fo.forward(sd, cfg, x, t, y, pos, mask, train=True, collect=streams)
    assert rel_l2(out[mask], torch.from_numpy(z["out_train"])[mask]) < tol
    if with_streams:
        ref_streams = torch.from_numpy(z["streams"])
        assert len(streams) == ref_streams.shape[0] == cfg["depth"] + 1
        for a, b in zip(streams, ref_streams):
            assert rel_l2(a[mask], b[mask]) < tol
    loss, _, grads = fo.loss_and_grads(sd, cfg, x, t, y, pos, mask, target)
    assert abs(float(loss) - float(z["loss"])) < 1e-5 * abs(float(z["loss"]))
    for k, g in grads.items():
        gn = float(z["gnorm." + k])
        assert abs(float(g.double().norm()) - gn) <= grad_tol * gn + 1e-12, k
        ref = torch.from_numpy(z["g." + k])
        got = sample_flat(g)
        assert float((got.double() - ref.double()).norm()) <= grad_tol * max(float(ref.double().norm()), 1e-3 * gn), k
    zz, tt, yy, pos_e, mask_e = synth.eval_batch(cfg, shapes[0][0], shapes[0][1], T)
    np.testing.assert_allclose(np.array([checksum(zz), float(yy.sum()), checksum(pos_e)]), z["eval_checksum"], rtol=1e-12)
    with torch.no_grad():
        assert rel_l2(fo.forward(sd, cfg, zz, tt, yy, pos_e, mask_e, train=False), z["out_eval"]) < tol
        assert rel_l2(fo.forward_with_cfg(sd, cfg, zz, tt, yy, pos_e, mask_e, 1.5), z["out_cfg"]) < tol


def test_fit_tiny_d64_matches_reference():
    _check_fit("fit_tiny_d64", TINY64, TINY_SHAPES, 64, 2e-6, True, 2e-5)


def test_fit_tiny_d72_matches_reference():
    _check_fit("fit_tiny_d72", TINY72, TINY_SHAPES, 64, 2e-6, True, 2e-5)


def test_fit_b2_mixed_matches_reference():
    _check_fit("fit_b2_mixed", fo.model_config("FiT-B/2"), [(32, 32), (20, 40), (16, 64)], 256, 5e-6, False, 5e-5)


def test_sample_loop_tiny_matches_reference():
    z = np.load(os.path.join(G, "sample_loop_tiny.npz"))
    sd = fo.random_state_dict(TINY64, seed=7)
    tb, tmap = do.create("10")
    zz = torch.from_numpy(z["z"]); yy = torch.from_numpy(z["y"])
    pos = torch.tensor(precompute_freqs_cis_2d(64, 8, 8, max_length=64), dtype=torch.float32)[None]
    mask = torch.ones(4, 64, dtype=torch.bool)
    noise = iter(torch.from_numpy(z["noise"]))
    model = lambda x, t, **kw: fo.forward_with_cfg(sd, TINY64, x, t, kw["y"], kw["pos"], kw["mask"], kw["cfg_scale"])
    kw = dict(y=yy, pos=pos, mask=mask, cfg_scale=1.5)
    trace = []
    do.sample_loop(tb, tmap, model, zz, lambda x: next(noise), "ddpm", model_kwargs=kw, trace=trace)
    ref = torch.from_numpy(z["ddpm.trace"])
    for i, (a, b) in enumerate(zip(trace, ref)):
        assert rel_l2(a, b) < 1e-4, i      # chaotic amplification of fp32 rounding over 10 steps
    noise = iter(torch.from_numpy(z["noise"]))
    out = do.sample_loop(tb, tmap, model, zz, lambda x: next(noise), "ddim", model_kwargs=kw)
    assert rel_l2(out, z["ddim.final"]) < 1e-4


def test_pad_helpers_match_reference():
    z = np.load(os.path.join(G, "pad_helpers.npz"))
    for (h, w) in [(32, 32), (20, 40), (16, 64), (40, 40)]:
        g = torch.Generator().manual_seed(h * 100 + w)
        lat = torch.randn(2, 4, h, w, generator=g)
        key = f"{h}x{w}"
        padded = fo.pad_latent(lat, 2, 32, 256)
        np.testing.assert_array_equal(padded.numpy(), z[key + ".padded"])
        pe, valid = fo.create_pos_embed(h, w, 2, 256, 64)
        assert valid == int(z[key + ".valid"]) and list(pe.shape) == list(z[key + ".pos_shape"])
        assert abs(checksum(pe) - float(z[key + ".pos_sum"])) < 1e-9 * float(z[key + ".pos_sum"])
        np.testing.assert_array_equal(fo.create_mask(valid, 256, 2).numpy(), z[key + ".mask"])
        if valid <= 256:
            np.testing.assert_array_equal(fo.unpad_latent(padded, valid, h, w, 2).numpy(), z[key + ".unpadded"])
            np.testing.assert_array_equal(fo.unpad_latent(padded, valid, h, w, 2).numpy(), lat.numpy())


def test_padded_tokens_do_not_influence_valid_tokens():
    # keys are masked, everything else is per-token
    cfg = TINY64
    sd = fo.random_state_dict(cfg, seed=3)
    x, t, y, pos, mask, _ = synth.mixed_batch(cfg, TINY_SHAPES, 64, seed=5, rope_dim=64)
    with torch.no_grad():
        a = fo.forward(sd, cfg, x, t, y, pos, mask)
        x2 = x.clone(); x2[~mask] = torch.randn_like(x2[~mask]) * 100
        b = fo.forward(sd, cfg, x2, t, y, pos, mask)
    assert torch.equal(a[mask], b[mask])


def test_masked_fit_collate_and_training_step_match_reference():
    """masked_FiT (row f4 + the masked leg of config 5): the restated subsample / collate reproduces, bit for bit, what the reference's
    own ImageNetLatentIterator._patchify + collate produced from the same items and RNG seeds (tokens in permuted order with their RoPE
    rows, padding for the short sample), and the oracle's training-step arithmetic on that batch matches the reference's loss and every
    parameter gradient (masked_FiT/train.py:40-60)."""
    from oracle.subsample_oracle import collate
    z = np.load(os.path.join(G, "masked_fit_tiny.npz"))
    toks, poss, masks, labels = synth.masked_items()
    T2 = int(z["number_of_tokens"])
    lat, pos, msk = collate(toks, poss, masks, T2, generator=torch.Generator().manual_seed(11))
    assert torch.equal(lat, torch.from_numpy(z["collate.latent"]))
    assert torch.equal(pos, torch.from_numpy(z["collate.pos"]))
    assert torch.equal(msk, torch.from_numpy(z["collate.mask"]))
    assert [int(v) for v in z["collate.label"].reshape(-1)] == labels
    assert int(msk.sum()) == 3 * 128 + 32 and not bool(msk[3, 32:].any())
    # the kept tokens are NOT a prefix of the image: the permuted RoPE rows are what tells the network where they came from
    assert not torch.equal(pos[0], poss[0][:T2])
    t, noise = synth.masked_step_noise(lat.shape)
    tb, _ = do.create("")
    x_t = do.q_sample(tb, lat, t, noise)
    assert rel_l2(x_t, z["x_t"]) < 1e-6
    cfg = synth.MASKED_CFG
    sd = fo.random_state_dict(cfg, seed=7)
    loss, out, grads = fo.loss_and_grads(sd, cfg, x_t, t, torch.tensor(labels), pos, msk, noise)
    assert rel_l2(out[msk], torch.from_numpy(z["out_train"])[msk]) < 2e-6
    assert abs(float(loss) - float(z["loss"])) < 1e-5 * abs(float(z["loss"]))
    for k, g in grads.items():
        assert rel_l2(g, z["g." + k]) < 2e-5, k
