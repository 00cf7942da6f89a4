"""Plain bf16 GEMM at training row counts (M = 8192): time per launch for N = 1152 / 3456 with K = 1152 / 3072, L2-warm back-to-back."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fit_b200 import ops
from fit_b200._lib import EPI_BIAS
M = int(os.environ.get("ROWS", 8192))
for K, N in ((1152, 1152), (3072, 1152), (1152, 3456), (1152, 6144)):
    a = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.03).bfloat16(); b = torch.randn(N, device="cuda")
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    f = lambda: ops.gemm(a, w, b, out, EPI_BIAS)
    for _ in range(5): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50): f()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 50
    print(f"M={M} K={K} N={N}: {ms*1e3:.1f} us = {2.0*M*N*K/ms/1e9:.0f} TF/s")
