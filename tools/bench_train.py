"""Time one training step (forward + backward of the custom autograd node + masked MSE) -- BASELINE config 4 shape per GPU:
FiT-XL/2, 32 samples x 256 tokens, bf16.  Prints ms, algorithmic TFLOP/s (3 x forward FLOPs) and the top kernels by device time."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from fit_b200.models.fit import FiT_models
from fit_b200.rope import precompute_freqs_cis_2d
from fit_b200.train import masked_mse_loss

name = os.environ.get("MODEL", "FiT-XL/2")
N, T = int(os.environ.get("BATCH", 32)), int(os.environ.get("TOK", 256))
valid = int(os.environ.get("VALID", T))
D, Hd, L, heads = bench.MODELS[name]
torch.manual_seed(0)
with torch.device("cuda"):
    m = FiT_models[name](precision=os.environ.get("PRECISION", "bf16"))
bench.randomize_zero_init(m)
m.train()
m.y_embedder.dropout_prob = 0.0
d = D // heads
pos = torch.zeros(N, T, d)
pos[:, :valid] = torch.tensor(precompute_freqs_cis_2d(d, 16, 16, max_length=256), dtype=torch.float32)[:valid]
pos = pos.cuda()
mask = torch.zeros(N, T, dtype=torch.bool, device="cuda"); mask[:, :valid] = True
x = torch.randn(N, T, 16, device="cuda"); tgt = torch.randn(N, T, 16, device="cuda")
t = torch.randint(0, 1000, (N,), device="cuda"); y = torch.randint(0, 1000, (N,), device="cuda")

def step():
    for p in m.parameters():
        p.grad = None
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e[0].record()
    out = m(x, t, y, pos, mask, train=True)
    loss = masked_mse_loss(out, tgt, mask)
    e[1].record()
    loss.backward()
    e[2].record()
    torch.cuda.synchronize()
    return e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2]), float(loss.detach())

for _ in range(2):
    step()
r = [step() for _ in range(int(os.environ.get("ITERS", 5)))]
f = sum(a for a, _, _ in r) / len(r); b = sum(bb for _, bb, _ in r) / len(r)
fl = 3 * N * bench.forward_flops(D, Hd, L, valid)
print(f"{name} train step N={N} T={T} valid={valid}: fwd {f:.2f} ms  bwd {b:.2f} ms  total {f + b:.2f} ms  = {fl / (f + b) / 1e9:.1f} TFLOP/s algorithmic; loss {r[-1][2]:.5f}")
if os.environ.get("PROFILE", "1") == "1":
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        step()
    agg = {}
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            k = ev.name[:90]
            a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += ev.device_time
    tot = sum(a[1] for a in agg.values())
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:22]:
        print(f"{a[1] / 1e3:9.3f} ms {a[0]:5d}x {100 * a[1] / tot:5.1f}%  {k}")
    print(f"{tot / 1e3:9.3f} ms total device time")
