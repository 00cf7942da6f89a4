"""ncu target for the per-step launch list: ONE un-captured FiT-XL/2 sampling step (the same C-ABI launches the CUDA graph replays) between
cudaProfilerStart / cudaProfilerStop, after model init and warm-up:

    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file launches.csv python tools/step_launches.py

so the list holds exactly the kernels of one step (no init kernels to skip).  tools/step_share.py turns it into per-kernel shares."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from fit_b200.diffusion import create_diffusion
from fit_b200.models.fit import FiT_models
from fit_b200.rope import precompute_freqs_cis_2d
from fit_b200.sampler import CFGSampler

name = os.environ.get("MODEL", "FiT-XL/2")
B = int(os.environ.get("BATCH", 256))
D, Hd, L, heads = bench.MODELS[name]
torch.manual_seed(0)
with torch.device("cuda"):
    m = FiT_models[name](precision="bf16")
bench.randomize_zero_init(m)
m.eval()
pos = torch.tensor(precompute_freqs_cis_2d(D // heads, 16, 16, max_length=256), dtype=torch.float32)[None]
s = CFGSampler(m, create_diffusion("250"), batch=B, latent_hw=(32, 32), cfg_scale=1.5, seed=1234, use_graph=False)
g = torch.Generator().manual_seed(1)
s.load(torch.randn(B, 4, 32, 32, generator=g), torch.randint(0, 1000, (B,), generator=g), pos)
s.run(3)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.profiler.start()
e0.record()
s.run(1)
e1.record()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print(f"one step: {e0.elapsed_time(e1):.3f} ms (un-captured launches), {s.kernels_per_step()} launches per step")
