"""ncu target: a few FiT-XL/2 blocks at the headline shape (R=512 rows x 256 tokens), forward only, so that
`ncu --set full -k regex:... -s 9 -c 9` lands on steady-state launches of each block kernel (7 launches per block).
PROF_TRAIN=1 adds one training step at the config-4 per-GPU shape (32 x 256) for the backward kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from fit_b200.models.fit import FiT_models
from fit_b200.rope import precompute_freqs_cis_2d

name = os.environ.get("MODEL", "FiT-XL/2")
R, T = int(os.environ.get("ROWS", 512)), 256
D, Hd, L, heads = bench.MODELS[name]
torch.manual_seed(0)
with torch.device("cuda"):
    m = FiT_models[name](precision="bf16")
bench.randomize_zero_init(m)
m.eval()
d = D // heads
pos = torch.tensor(precompute_freqs_cis_2d(d, 16, 16, max_length=256), dtype=torch.float32)[None].cuda()
mask = torch.ones(R, T, dtype=torch.bool, device="cuda")
x = torch.randn(R, 4, 32, 32, device="cuda")
t = torch.full((R,), 500, dtype=torch.int64, device="cuda")
y = torch.randint(0, 1000, (R,), device="cuda")
eng = m.engine
W = eng.weights()
ws = eng.workspace(R, T, x.device)
with torch.no_grad():
    m(x, t, y, pos, mask, train=False)          # conditioning + every buffer initialised
    torch.cuda.synchronize()
    for i in range(int(os.environ.get("BLOCKS", 4))):
        eng.block(i, ws, W, pos, mask)
    torch.cuda.synchronize()
if os.environ.get("PROF_TRAIN", "0") == "1":
    from fit_b200.train import masked_mse_loss
    N = 32
    m.train(); m.y_embedder.dropout_prob = 0.0
    xt = torch.randn(N, T, 16, device="cuda"); tgt = torch.randn(N, T, 16, device="cuda")
    posb = pos.expand(N, T, d).contiguous()
    out = m(xt, t[:N], y[:N], posb, mask[:N], train=True)
    masked_mse_loss(out, tgt, mask[:N]).backward()
    torch.cuda.synchronize()
print("prof_kernels done")
