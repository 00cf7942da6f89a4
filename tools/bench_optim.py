"""Time the fused AdamW + EMA step on all FiT-XL/2 parameters (674.9 M fp32 elements, 36 B / element algorithmic) against torch.optim.AdamW
(foreach) + the reference's foreach EMA update on the same GPU."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from fit_b200.models.fit import FiT_models
from fit_b200.optim import FusedAdamWEMA

with torch.device("cuda"):
    m = FiT_models[os.environ.get("MODEL", "FiT-XL/2")](precision="bf16")
params = [p for p in m.parameters()]
n = sum(p.numel() for p in params)
for p in params:
    p.grad = torch.randn_like(p) * 1e-3
opt = FusedAdamWEMA(params, lr=1e-4, ema_decay=0.9999)


def timed(fn, k=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(k):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / k


t = timed(opt.step)
peaks = bench.load_peaks()
print(f"fused AdamW+EMA: {n / 1e6:.1f} M elements, {t:.3f} ms/step = {n * 36 / t / 1e6:.0f} GB/s algorithmic ({n * 36 / t / 1e6 / peaks['hbm_gbs']:.2f} of {peaks['hbm_gbs']:.0f} GB/s)")
del opt
ref = torch.optim.AdamW(params, lr=1e-4, weight_decay=0, foreach=True)
ema = [p.detach().clone() for p in params]


def ref_step():
    ref.step()
    torch._foreach_mul_(ema, 0.9999)
    torch._foreach_add_(ema, [p.detach() for p in params], alpha=1e-4)


with torch.no_grad():
    t2 = timed(ref_step)
print(f"torch AdamW(foreach) + foreach EMA (the reference's optimizer path, FiT/train.py:215-217, ema.py:149-154): {t2:.3f} ms/step -> {t2 / t:.2f}x")
