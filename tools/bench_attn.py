"""Time the attention kernel alone at the headline shape (R=512 rows x 256 tokens, FiT-XL/2 heads) with CUDA events, L2 flushed."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fit_b200 import ops
from fit_b200.rope import precompute_freqs_cis_2d

R, T, h, d = int(os.environ.get("ROWS", 512)), int(os.environ.get("TOK", 256)), 16, int(os.environ.get("HD", 72))
valid = int(os.environ.get("VALID", T))
qkv = torch.randn(R * T, 3 * h * d, device="cuda").bfloat16()
pos = torch.tensor(precompute_freqs_cis_2d(d, 16, T // 16, max_length=256), dtype=torch.float32)[None].cuda()
mask = torch.zeros(R, T, dtype=torch.bool, device="cuda"); mask[:, :valid] = True
kv, _ = ops.mask_to_kvlen(mask)
out = torch.empty(R * T, h * d, device="cuda", dtype=torch.bfloat16)
scratch = torch.empty(64 * 1024 * 1024, device="cuda")
PRE = 1   # bf16 operands always carry the RoPE (the un-rotated bf16 kernels were removed in round 2)
run = lambda: ops.attention(qkv, None, kv, mask, R, T, h, d, out, prerotated=True)
for _ in range(3):
    run()
torch.cuda.synchronize()
tot = 0.0
N = 10
for _ in range(N):
    scratch.fill_(1.0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); run(); e1.record()
    torch.cuda.synchronize(); tot += e0.elapsed_time(e1)
ms = tot / N
fl = 4.0 * R * valid * valid * h * d
by = (R * valid * 3 * h * d + R * valid * h * d) * 2
print(f"attention pre={PRE} R={R} T={T} valid={valid} h={h} d={d}: {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s  {by / ms / 1e6:.1f} GB/s (algorithmic)")
