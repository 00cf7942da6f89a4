"""Timings of the BASELINE.json configs that are not the bench line (configs 2 and 5): parity-test cases, measured once for DESIGN.md.
  config 2  FiT-B/2 mixed-resolution forward+backward (tools/bench_train.py MODEL=FiT-B/2 VALID=200)
  config 5  masked_FiT: XL/2 train step with 50 % of the tokens kept (VALID=128) -- run via tools/bench_train.py -- and
            resolution-extrapolation sampling at 320x320 (400 tokens) and 1024x1024 (4096 tokens): this script."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from fit_b200.diffusion import create_diffusion
from fit_b200.models.fit import FiT_models
from fit_b200.rope import create_pos_embed
from fit_b200.sampler import CFGSampler

name = "FiT-XL/2"
D, Hd, L, heads = bench.MODELS[name]
torch.manual_seed(0)
with torch.device("cuda"):
    m = FiT_models[name](precision="bf16")
bench.randomize_zero_init(m)
m.eval()
diff = create_diffusion("250")
for side, B in ((320, 64), (1024, 4)):
    hl = side // 8
    table, valid_t = create_pos_embed(hl, hl, 2, 256, D // heads)       # VisionNTK scaling beyond 256 tokens (pos_embed.py:51-75)
    pos = table[:, :valid_t].contiguous()
    s = CFGSampler(m, diff, batch=B, latent_hw=(hl, hl), cfg_scale=1.5, seed=0)
    z = torch.randn(B, 4, hl, hl)
    y = torch.randint(0, 1000, (B,))
    s.load(z, y, pos)
    s.run(2)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    K = 5
    e0.record(); s.run(K); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / K
    fl = 2 * B * bench.forward_flops(D, Hd, L, valid_t)
    assert bool(torch.isfinite(s.x).all())
    print(f"{name} sampling at {side}x{side} ({valid_t} tokens), batch {B} ({2 * B} rows): {ms:.2f} ms/step = {B / (250 * ms * 1e-3):.3f} img/s "
          f"(250-step DDPM, CFG), {fl / ms / 1e9:.0f} TFLOP/s algorithmic")
    del s
