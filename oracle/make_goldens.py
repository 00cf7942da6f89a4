"""Generate tests/golden/*.npz by running the UNMODIFIED reference (jlnk03/FiT) on CPU.

Run in the build container only (needs /root/reference):

    python oracle/make_goldens.py

The reference is imported from /root/reference/FiT with stub modules for the packages it
imports but the hot path never uses (timm in models/dit.py:16; lightning / diffusers / wandb /
torchvision pieces in train.py:5-16).  Weights and inputs are regenerated at test time from the
seeds recorded here (oracle.fit_oracle.random_state_dict -- same torch build on both machines),
so the fixtures hold only expected OUTPUTS plus checksums of the inputs.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference/FiT"
OUT = os.path.join(ROOT, "tests", "golden")


def _stub(name, **attrs):
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


def import_reference():
    class _Placeholder(torch.nn.Module):
        def __init__(self, *a, **k):
            super().__init__()

    _stub("timm"); _stub("timm.models")
    _stub("timm.models.vision_transformer", PatchEmbed=_Placeholder, Attention=_Placeholder, Mlp=_Placeholder)

    class _LM(torch.nn.Module):
        device = torch.device("cpu")
        def save_hyperparameters(self, *a, **k): pass
        def log(self, *a, **k): pass
    _stub("lightning", LightningModule=_LM, Trainer=object, seed_everything=lambda *a, **k: None)
    _stub("lightning.pytorch"); _stub("lightning.pytorch.callbacks", ModelCheckpoint=object, Callback=object)
    _stub("lightning.pytorch.loggers", WandbLogger=object)
    _stub("lightning.pytorch.profilers", AdvancedProfiler=object)
    _stub("lightning.pytorch.utilities"); _stub("lightning.pytorch.utilities.exceptions", MisconfigurationException=Exception)
    _stub("lightning.pytorch.utilities.rank_zero", rank_zero_info=print)
    _stub("diffusers", DDIMScheduler=object); _stub("diffusers.models", AutoencoderKL=object)
    _stub("torchvision"); _stub("torchvision.utils", save_image=None)
    _stub("ema", EMA=object)
    sys.path.insert(0, REF)
    from models.fit import FiT, FiT_models  # noqa
    from diffusion import create_diffusion  # noqa
    from preprocess.pos_embed import precompute_freqs_cis_2d  # noqa
    return FiT, FiT_models, create_diffusion, precompute_freqs_cis_2d


def checksum(t: torch.Tensor) -> float:
    return float(t.double().abs().sum())


def sample_flat(t: torch.Tensor, n: int = 2048) -> np.ndarray:
    f = t.reshape(-1)
    step = max(1, f.numel() // n)
    return f[::step][:n].numpy().copy()


from oracle.synth import mixed_batch, eval_batch  # noqa: E402


def gen_fit(name, cfg, shapes, T, FiT, precompute, full_grads=False, blocks=True):
    from oracle import fit_oracle as fo
    sd = fo.random_state_dict(cfg, seed=7)
    ref = FiT(**{k: cfg[k] for k in ("patch_size", "in_channels", "hidden_size", "depth", "num_heads",
                                       "mlp_ratio", "class_dropout_prob", "num_classes", "learn_sigma")})
    ref.load_state_dict(sd)
    ref.eval()  # label dropout off (fit.py:504 keys on self.training)
    d = cfg["hidden_size"] // cfg["num_heads"]
    x, t, y, pos, mask, target = mixed_batch(cfg, shapes, T, seed=11, rope_dim=d, precompute=precompute)
    out = {}
    out["in_checksum"] = np.array([checksum(x), checksum(pos), float(t.sum()), float(y.sum()), checksum(target),
                                   sum(checksum(v) for v in sd.values())])
    # forward (train=True: tokens in, tokens out) with per-block hooks
    streams = []
    hooks = []
    if blocks:
        hooks.append(ref.x_embedder.register_forward_hook(lambda m, i, o: streams.append(o.detach().clone())))
        for b in ref.blocks:
            hooks.append(b.register_forward_hook(lambda m, i, o: streams.append(o.detach().clone())))
    for prm in ref.parameters():
        prm.grad = None
    y_out = ref(x, t, y, pos, mask, train=True)
    for h in hooks:
        h.remove()
    loss = torch.nn.functional.mse_loss(y_out[mask], target[mask])   # FiT/train.py:63-66
    loss.backward()
    out["out_train"] = y_out.detach().numpy()
    out["loss"] = np.array(float(loss))
    if blocks:
        out["streams"] = torch.stack(streams).numpy()
    for k, prm in ref.named_parameters():
        g = prm.grad
        out["gnorm." + k] = np.array(float(g.double().norm()))
        out["g." + k] = g.numpy().copy() if full_grads else sample_flat(g)
    # eval path: NCHW in/out + CFG, single resolution (first shape), all-ones mask
    h, w = shapes[0]
    zz, tt, yy, pos_e, mask_e = eval_batch(cfg, h, w, T, precompute=precompute)
    with torch.no_grad():
        out["out_eval"] = ref(zz, tt, yy, pos_e, mask_e, train=False).numpy()
        out["out_cfg"] = ref.forward_with_cfg(zz, tt, yy, pos_e, mask_e, 1.5).numpy()
    out["eval_checksum"] = np.array([checksum(zz), float(yy.sum()), checksum(pos_e)])
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(name, "loss", float(loss), "out absmax", float(y_out.abs().max()))
    return ref, sd


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_grad_enabled(True)
    FiT, FiT_models, create_diffusion, precompute = import_reference()
    from oracle import fit_oracle as fo

    # ---- 1. RoPE tables (pos_embed.py:51-75)
    rope = {}
    for (dim, nh, nw, ml) in [(64, 16, 16, None), (64, 16, 16, 256), (72, 16, 16, 256), (64, 10, 20, 256),
                              (64, 8, 32, 256), (72, 20, 20, 256), (72, 64, 64, 256), (64, 4, 4, 64), (72, 5, 10, 64)]:
        tab = precompute(dim, nh, nw, max_length=ml)
        key = f"d{dim}_h{nh}_w{nw}_ml{ml}"
        if tab.size <= 40000:
            rope[key] = tab
        else:
            rope[key + "_sample"] = tab.reshape(-1)[:: tab.size // 4096][:4096].copy()
            rope[key + "_sum"] = np.array(np.abs(tab).sum())
    np.savez_compressed(os.path.join(OUT, "rope_tables.npz"), **rope)

    # ---- 2. diffusion tables (gaussian_diffusion.py:165-213, respace.py:65-92)
    tabs = {}
    for rs in ["10", "250", "ddim25", ""]:
        d = create_diffusion(rs)
        tag = rs or "full"
        tabs[tag + ".timestep_map"] = np.array(d.timestep_map)
        for f in ["betas", "alphas_cumprod", "alphas_cumprod_prev", "sqrt_alphas_cumprod",
                  "sqrt_one_minus_alphas_cumprod", "sqrt_recip_alphas_cumprod", "sqrt_recipm1_alphas_cumprod",
                  "posterior_variance", "posterior_log_variance_clipped", "posterior_mean_coef1",
                  "posterior_mean_coef2"]:
            tabs[tag + "." + f] = getattr(d, f)
    np.savez_compressed(os.path.join(OUT, "diffusion_tables.npz"), **tabs)

    # ---- 3. single diffusion steps with a fixed "model" output
    steps = {}
    d = create_diffusion("10")
    g = torch.Generator().manual_seed(5)
    x = torch.randn(6, 4, 8, 8, generator=g)
    eps = torch.randn(6, 4, 8, 8, generator=g)
    t = torch.tensor([9, 5, 0, 1, 0, 7])
    steps["x"], steps["eps"], steps["t"] = x.numpy(), eps.numpy(), t.numpy()
    model = lambda x_, t_, **kw: eps
    torch.manual_seed(99); noise = torch.randn_like(x); steps["noise"] = noise.numpy()
    torch.manual_seed(99); o = d.p_sample(model, x, t, clip_denoised=False)
    steps["p_sample.sample"], steps["p_sample.pred_xstart"] = o["sample"].numpy(), o["pred_xstart"].numpy()
    pm = d.p_mean_variance(model, x, t, clip_denoised=False)
    for k in ("mean", "variance", "log_variance"):
        steps["pmv." + k] = pm[k].numpy()
    for eta in (0.0, 0.5):
        torch.manual_seed(99); o = d.ddim_sample(model, x, t, clip_denoised=False, eta=eta)
        steps[f"ddim{eta}.sample"] = o["sample"].numpy()
    steps["q_sample"] = d.q_sample(x, t, noise=noise).numpy()
    np.savez_compressed(os.path.join(OUT, "diffusion_steps.npz"), **steps)

    # ---- 4. tiny FiT models at head_dim 64 and 72 (full grads), mixed resolution, T = 64
    tiny64 = dict(patch_size=2, in_channels=4, hidden_size=128, depth=2, num_heads=2, mlp_ratio=4.0,
                  class_dropout_prob=0.1, num_classes=10, learn_sigma=False)
    tiny72 = dict(tiny64, hidden_size=144)
    shapes = [(16, 16), (10, 20), (8, 32), (12, 12)]
    ref64, sd64 = gen_fit("fit_tiny_d64", tiny64, shapes, 64, FiT, precompute)
    gen_fit("fit_tiny_d72", tiny72, shapes, 64, FiT, precompute)

    # ---- 5. real widths: FiT-B/2 mixed-res (BASELINE config 2 shape) and FiT-XL/2 (2 rows), outputs + sampled grads
    b2 = fo.model_config("FiT-B/2")
    gen_fit("fit_b2_mixed", b2, [(32, 32), (20, 40), (16, 64)], 256, FiT, precompute, blocks=False)
    xl2 = fo.model_config("FiT-XL/2")
    gen_fit("fit_xl2_small", xl2, [(32, 32), (20, 40)], 256, FiT, precompute, blocks=False)

    # ---- 6. 10-step DDPM + DDIM sampling loops with CFG on the tiny d64 model (config-1 shape, scaled down)
    loop = {}
    d = create_diffusion("10")
    g = torch.Generator().manual_seed(21)
    B = 2
    z = torch.randn(B, 4, 16, 16, generator=g)
    zz = torch.cat([z, z], 0)
    yy = torch.cat([torch.randint(0, 10, (B,), generator=g), torch.full((B,), 10)])
    pos = torch.tensor(precompute(64, 8, 8, max_length=64), dtype=torch.float32)[None]
    mask = torch.ones(2 * B, 64, dtype=torch.bool)
    kw = dict(y=yy, pos=pos, mask=mask, cfg_scale=1.5)
    torch.manual_seed(123)
    trace = [o["sample"].numpy() for o in d.p_sample_loop_progressive(
        ref64.forward_with_cfg, zz.shape, zz, clip_denoised=False, model_kwargs=kw, device="cpu")]
    loop["ddpm.trace"] = np.stack(trace)
    torch.manual_seed(123)
    loop["ddim.final"] = d.ddim_sample_loop(ref64.forward_with_cfg, zz.shape, zz, clip_denoised=False,
                                            model_kwargs=kw, device="cpu").numpy()
    torch.manual_seed(123)
    loop["noise"] = np.stack([torch.randn_like(zz).numpy() for _ in range(10)])
    loop["z"] = zz.numpy(); loop["y"] = yy.numpy()
    np.savez_compressed(os.path.join(OUT, "sample_loop_tiny.npz"), **loop)

    # ---- 7. padding helpers of the sampling script (FiT/train.py:97-166)
    import train as ref_train
    mod = ref_train.FiTModule.__new__(ref_train.FiTModule)
    torch.nn.Module.__init__(mod)
    hp = {}
    for (h, w) in [(32, 32), (20, 40), (16, 64), (40, 40)]:
        g = torch.Generator().manual_seed(h * 100 + w)
        lat = torch.randn(2, 4, h, w, generator=g)
        padded = mod._pad_latent(lat, 2, 32, 256)
        pe, valid = mod._create_pos_embed(h, w, 2, 256, 64, "rotate")
        mk = mod._create_mask(valid, 256, 2)
        key = f"{h}x{w}"
        hp[key + ".padded"] = padded.numpy(); hp[key + ".pos_sum"] = np.array(float(pe.double().abs().sum()))
        hp[key + ".pos_shape"] = np.array(pe.shape); hp[key + ".valid"] = np.array(valid)
        hp[key + ".mask"] = mk.numpy()
        if valid <= 256:
            hp[key + ".unpadded"] = mod._unpad_latent(padded, valid, h, w, 2).numpy()
    np.savez_compressed(os.path.join(OUT, "pad_helpers.npz"), **hp)
    gen_masked(FiT, create_diffusion)
    print("goldens written to", OUT)


from oracle.synth import MASKED_CFG, MASKED_SHAPES, masked_inputs  # noqa: E402


def gen_masked(FiT, create_diffusion):
    """masked_FiT: the reference's OWN ImageNetLatentIterator._patchify / subsample / collate (masked_FiT/preprocess/iterators.py:128-169,
    called unbound: the class constructor needs dataset files) followed by its training-step arithmetic (masked_FiT/train.py:40-60) on
    the reference model.  Pins row f4 (token sub-sampling in permuted order with the matching RoPE rows) and the masked leg of config 5."""
    import importlib.util
    import random
    if "torchvision.transforms" not in sys.modules:
        _stub("torchvision.transforms")
    spec = importlib.util.spec_from_file_location("masked_iterators", "/root/reference/masked_FiT/preprocess/iterators.py")
    mi = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mi)
    it = mi.ImageNetLatentIterator.__new__(mi.ImageNetLatentIterator)
    it.patch_size, it.C, it.embed_dim, it.embed_method = 2, 4, 72, "rotate"
    lats, labels = masked_inputs()
    batch = []
    for lat, lab in zip(lats, labels):
        tok, pos = it._patchify(lat)                                   # iterators.py:128-140
        batch.append((torch.tensor(tok), [lab], torch.tensor(pos), torch.ones(tok.shape[0], dtype=torch.bool)))
    out = {}
    seed_py = next(s for s in range(100) if random.Random(s).choice([32, 64, 96, 128, 192, 256]) == 128)
    random.seed(seed_py)                                               # collate draws number_of_tokens with random.choice (:158)
    torch.manual_seed(11)                                              # subsample draws torch.randperm from the global generator (:144)
    latent, label, pos, mask = it.collate(batch)
    assert latent.shape[1] == 128
    out["collate.latent"], out["collate.pos"], out["collate.mask"] = latent.numpy(), pos.numpy(), mask.numpy()
    out["collate.label"] = label.numpy()
    out["number_of_tokens"] = np.array(128)
    # training step (masked_FiT/train.py:40-60); DDIMScheduler.add_noise restated by the reference's own q_sample (same linear schedule)
    from oracle import fit_oracle as fo
    sd = fo.random_state_dict(MASKED_CFG, seed=7)
    ref = FiT(**MASKED_CFG)
    ref.load_state_dict(sd)
    ref.eval()
    g = torch.Generator().manual_seed(17)
    t = torch.randint(0, 1000, (latent.shape[0],), generator=g)
    noise = torch.randn(latent.shape, generator=g)
    x_t = create_diffusion("").q_sample(latent, t, noise=noise)
    y = label.to(torch.int64).reshape(-1)
    model_output = ref(x_t, t=t, y=y, pos=pos, mask=mask)
    loss = torch.nn.functional.mse_loss(model_output[mask], noise[mask])
    loss.backward()
    out["x_t"], out["out_train"], out["loss"] = x_t.numpy(), model_output.detach().numpy(), np.array(float(loss))
    for k, prm in ref.named_parameters():
        out["gnorm." + k] = np.array(float(prm.grad.double().norm()))
        out["g." + k] = prm.grad.numpy().copy()
    np.savez_compressed(os.path.join(OUT, "masked_fit_tiny.npz"), **out)
    print("masked_fit_tiny loss", float(loss), "valid tokens", int(mask.sum()))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "masked":   # only the masked_FiT fixture (the others are unchanged)
        FiT_, _, create_diffusion_, _ = import_reference()
        gen_masked(FiT_, create_diffusion_)
    else:
        main()
