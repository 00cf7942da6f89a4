"""The GEMM shapes of one FiT-XL/2 training block at 8192 token rows (fwd, dgrad, wgrad), ours vs cuBLAS (torch.matmul), back to back."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fit_b200 import ops
from fit_b200._lib import EPI_BIAS
M, D, H = int(os.environ.get("ROWS", 8192)), 1152, 3072
bf = torch.bfloat16
def t(fn, n=30):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
tot_o = tot_c = 0.0
for name, n_out, k_in in (("qkv", 3 * D, D), ("proj", D, D), ("fc1", 2 * H, D), ("fc2", D, H)):
    x = torch.randn(M, k_in, device="cuda").to(bf); w = (torch.randn(n_out, k_in, device="cuda") * 0.03).to(bf)
    dy = torch.randn(M, n_out, device="cuda").to(bf)
    y = torch.empty(M, n_out, dtype=bf, device="cuda"); dx = torch.empty(M, k_in, dtype=bf, device="cuda"); dw = torch.empty(n_out, k_in, device="cuda")
    dwb = torch.empty(n_out, k_in, dtype=bf, device="cuda")
    rows = [("fwd", lambda: ops.gemm(x, w, None, y, EPI_BIAS), lambda: torch.matmul(x, w.t(), out=y)),
            ("dgrad", lambda: ops.gemm(dy, w, None, dx, EPI_BIAS, b_mn=True), lambda: torch.matmul(dy, w, out=dx)),
            ("wgrad", lambda: ops.gemm(dy, x, None, dw, EPI_BIAS, a_mn=True, b_mn=True), lambda: torch.matmul(dy.t(), x, out=dwb))]
    for kind, ours, cub in rows:
        a, b = t(ours), t(cub)
        tot_o += a; tot_c += b
        fl = 2.0 * M * n_out * k_in
        print(f"{name:5s} {kind:5s}: ours {a:7.1f} us = {fl/a/1e6:6.0f} TF/s | cuBLAS {b:7.1f} us = {fl/b/1e6:6.0f} TF/s | {a/b:.3f}")
print(f"block total: ours {tot_o:.0f} us, cuBLAS {tot_c:.0f} us (cuBLAS wgrad writes bf16, ours fp32)")
