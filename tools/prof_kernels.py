"""Launch each hot kernel of one FiT-XL/2 block once at the headline shape (R=512 rows x 256 tokens) -- the command ncu wraps."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fit_b200.models.fit import FiT
from fit_b200.rope import precompute_freqs_cis_2d
import bench

R = int(os.environ.get("PROF_ROWS", "512"))
name = os.environ.get("PROF_MODEL", "FiT-XL/2")
torch.manual_seed(0)
with torch.device("cuda"):
    cfgs = {"FiT-XL/2": dict(hidden_size=1152, num_heads=16), "FiT-B/2": dict(hidden_size=768, num_heads=12)}
    m = FiT(depth=2, patch_size=2, **cfgs[name])
bench.randomize_zero_init(m)
m = m.eval()
dh = m.hidden_size // m.num_heads
pos = torch.tensor(precompute_freqs_cis_2d(dh, 16, 16, max_length=256), dtype=torch.float32)[None].cuda()
x = torch.randn(R // 2, 4, 32, 32, device="cuda")
t = torch.full((R,), 500, device="cuda")
y = torch.randint(0, 1000, (R,), device="cuda")
mask = torch.ones(R, 256, dtype=torch.bool, device="cuda")
with torch.no_grad():
    for _ in range(int(os.environ.get("PROF_ITERS", "2"))):
        out = m.engine.forward(x, t, y, pos, mask, train=False, rows=R)
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()))
