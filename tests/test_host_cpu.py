"""CPU-side tests of the host logic: the C ABI loads and exports what the header declares, the diffusion tables,
the RoPE table generator and the padding helpers reproduce the reference's fixtures, and the product path refuses
to run without CUDA (no fallback).  No kernel is launched here."""
import os
import re

import numpy as np
import pytest
import torch

from helpers import GOLDEN, TINY64, golden


def test_library_exports_every_declared_symbol():
    from fit_b200 import _lib
    names = _lib.declared_symbols()
    assert len(names) >= 19 and "fitb200_gemm_bf16_tc" in names and "fitb200_attn_fwd" in names
    lib = _lib.lib()   # raises if the .so is missing
    protos = _lib.declared_prototypes()
    for n in names:
        assert hasattr(lib, n), n
        assert n in protos, n
    assert lib.fitb200_version() >= 100


def test_header_cites_reference_lines():
    with open(os.path.join(os.path.dirname(GOLDEN), "..", "include", "fit_b200.h")) as fh:
        src = fh.read()
    assert len(re.findall(r"(fit|gaussian_diffusion|dit|respace)\.py:\d+", src)) >= 15


def test_gemm_args_struct_matches_header_layout(tmp_path):
    """Every field of the ctypes mirror sits at the offset the C compiler gives fitb200_gemm_args (gcc on include/fit_b200.h)."""
    import ctypes
    import subprocess
    from fit_b200._lib import GemmArgs, HEADER_PATH
    names = [f[0] for f in GemmArgs._fields_]
    src = tmp_path / "off.c"
    src.write_text('#include <stddef.h>\n#include <stdio.h>\n#include "%s"\nint main(void) {\n' % HEADER_PATH +
                   "".join('  printf("%s %%zu\\n", offsetof(fitb200_gemm_args, %s));\n' % (n, n) for n in names) +
                   '  printf("sizeof %zu\\n", sizeof(fitb200_gemm_args));\n  return 0;\n}\n')
    exe = tmp_path / "off"
    subprocess.run(["gcc", str(src), "-o", str(exe)], check=True)
    lines = dict(l.split() for l in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    for n in names:
        assert getattr(GemmArgs, n).offset == int(lines[n]), n
    assert ctypes.sizeof(GemmArgs) == int(lines["sizeof"])
    # the round-1 fields keep their offsets (new fields are appended)
    assert GemmArgs.M.offset == 0 and GemmArgs.A.offset == 16 and GemmArgs.rope_freqs.offset == 144 and GemmArgs.b_mn.offset == 172


@pytest.mark.parametrize("rs", ["10", "250", "ddim25", ""])
def test_diffusion_tables_match_reference(rs):
    from fit_b200.diffusion import create_diffusion
    z = golden("diffusion_tables")
    d = create_diffusion(rs)
    tag = rs or "full"
    np.testing.assert_array_equal(np.array(d.timestep_map), z[tag + ".timestep_map"])
    for f in ["betas", "alphas_cumprod", "alphas_cumprod_prev", "sqrt_alphas_cumprod", "sqrt_one_minus_alphas_cumprod",
              "sqrt_recip_alphas_cumprod", "sqrt_recipm1_alphas_cumprod", "posterior_variance", "posterior_log_variance_clipped",
              "posterior_mean_coef1", "posterior_mean_coef2"]:
        np.testing.assert_array_equal(getattr(d, f), z[tag + "." + f], err_msg=f)
    assert d.num_timesteps == len(d.timestep_map)


def test_create_diffusion_options():
    from fit_b200.diffusion import create_diffusion, gaussian_diffusion as gd
    d = create_diffusion("50", sigma_small=True, predict_xstart=True)
    assert d.model_var_type == gd.ModelVarType.FIXED_SMALL and d.model_mean_type == gd.ModelMeanType.START_X
    with pytest.raises(NotImplementedError):
        create_diffusion("50", learn_sigma=True)
    with pytest.raises(ValueError):
        create_diffusion("2000")


def test_rope_tables_match_reference():
    from fit_b200.rope import precompute_freqs_cis_2d
    z = golden("rope_tables")
    for key in z.files:
        parts = key.split("_")
        dim, nh, nw = int(parts[0][1:]), int(parts[1][1:]), int(parts[2][1:])
        ml = None if parts[3] == "mlNone" else int(parts[3][2:])
        tab = precompute_freqs_cis_2d(dim, nh, nw, max_length=ml)
        if key.endswith("_sample"):
            np.testing.assert_array_equal(tab.reshape(-1)[:: tab.size // 4096][:4096], z[key])
        elif key.endswith("_sum"):
            assert np.abs(tab).sum() == z[key]
        else:
            np.testing.assert_array_equal(tab, z[key])


def test_padding_helpers_match_reference():
    from fit_b200 import rope
    z = golden("pad_helpers")
    for (h, w) in [(32, 32), (20, 40), (16, 64), (40, 40)]:
        g = torch.Generator().manual_seed(h * 100 + w)
        lat = torch.randn(2, 4, h, w, generator=g)
        key = f"{h}x{w}"
        padded = rope.pad_latent(lat, 2, 32, 256)
        np.testing.assert_array_equal(padded.numpy(), z[key + ".padded"])
        pe, valid = rope.create_pos_embed(h, w, 2, 256, 64)
        assert valid == int(z[key + ".valid"]) and list(pe.shape) == list(z[key + ".pos_shape"])
        assert float(pe.double().abs().sum()) == float(z[key + ".pos_sum"])
        np.testing.assert_array_equal(rope.create_mask(valid, 256, 2).numpy(), z[key + ".mask"])
        if valid <= 256:
            np.testing.assert_array_equal(rope.unpad_latent(padded, valid, h, w, 2).numpy(), z[key + ".unpadded"])


def test_model_interface_matches_reference_names():
    from fit_b200.models.fit import FiT, FiT_models
    from oracle import fit_oracle as fo
    assert sorted(FiT_models) == sorted(f"FiT-{s}/{p}" for s in ("S", "B", "L", "XL") for p in (2, 4, 8))
    m = FiT(**TINY64)
    want = fo.param_shapes(TINY64)
    got = {k: tuple(v.shape) for k, v in m.state_dict().items()}
    assert got == want
    # adaLN-Zero init: modulation and output layers are zero (fit.py:421-430)
    assert float(m.blocks[0].adaLN_modulation[1].weight.abs().sum()) == 0.0
    assert float(m.final_layer.linear.weight.abs().sum()) == 0.0
    x = torch.randn(2, 4, 8, 8)
    assert torch.equal(m.unpatchify(m.patchify(x), 8, 8), x)
    assert torch.equal(m.patchify(x), fo.patchify(x, 2))
    with pytest.raises(TypeError):
        FiT(**dict(TINY64, ffn="mlp"))
    with pytest.raises(AssertionError):
        FiT(**TINY64, block_kwargs={"enable_flash_attention": True})


def test_product_path_refuses_cpu_tensors():
    from fit_b200.models.fit import FiT
    from oracle import synth
    m = FiT(**TINY64).eval()
    x, t, y, pos, mask, _ = synth.mixed_batch(TINY64, [(16, 16), (8, 8)], 64, seed=1, rope_dim=64)
    with torch.no_grad(), pytest.raises(RuntimeError, match="CUDA"):
        m(x, t, y, pos, mask)


def test_product_code_never_imports_the_oracle():
    pkg = os.path.join(os.path.dirname(GOLDEN), "..", "fit_b200")
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                with open(os.path.join(root, f)) as fh:
                    src = fh.read()
                assert "import oracle" not in src and "from oracle" not in src, f
