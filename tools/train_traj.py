"""Loss trajectory of the XL/2 training step on one fixed synthetic batch: eager optimizer steps vs CapturedTrainStep (debug / evidence)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from fit_b200.models.fit import FiT_models
from fit_b200.optim import FusedAdamWEMA
from fit_b200.rope import precompute_freqs_cis_2d
from fit_b200.train import CapturedTrainStep, attach_optimizer, masked_mse_loss

name = os.environ.get("MODEL", "FiT-B/2")
D, Hd, L, heads = bench.MODELS[name]
T, d, N = 256, D // heads, 16
res = {}
for mode in ("eager", "eager_plain", "graph"):
    torch.manual_seed(0)
    with torch.device("cuda"):
        m = FiT_models[name](precision="bf16")
    bench.randomize_zero_init(m)
    m.train(); m.y_embedder.dropout_prob = 0.0
    g = torch.Generator().manual_seed(5)
    x = torch.randn(N, T, 16, generator=g).cuda(); tgt = torch.randn(N, T, 16, generator=g).cuda()
    t = torch.randint(0, 1000, (N,), generator=g).cuda(); y = torch.randint(0, 1000, (N,), generator=g).cuda()
    pos = torch.tensor(precompute_freqs_cis_2d(d, 16, 16, max_length=256), dtype=torch.float32)[None].expand(N, T, d).contiguous().cuda()
    mask = torch.ones(N, T, dtype=torch.bool, device="cuda")
    opt = FusedAdamWEMA(m.parameters(), lr=1e-4, ema_decay=0.9999)
    losses = []
    if mode == "graph":
        step = CapturedTrainStep(m, opt)
        for k in range(8):
            losses.append(float(step(x, t, y, pos, mask, tgt)))
    else:
        if mode == "eager":
            attach_optimizer(m, opt)
        for k in range(8):
            for p in m.parameters():
                p.grad = None
            loss = masked_mse_loss(m(x, t, y, pos, mask, train=True), tgt, mask)
            loss.backward(); opt.step()
            losses.append(float(loss))
    res[mode] = losses
    print(mode, " ".join(f"{v:.5f}" for v in losses))
    del m, opt
    torch.cuda.empty_cache()
