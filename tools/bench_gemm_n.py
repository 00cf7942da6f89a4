"""Plain bf16 GEMM rate vs N (tile quantisation of the 256-wide pair tiles): M = 131072, K from argv."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fit_b200 import ops
from fit_b200._lib import EPI_BIAS
M = 131072
K = int(sys.argv[1]) if len(sys.argv) > 1 else 1152
a = torch.randn(M, K, device="cuda").bfloat16()
for N in (1024, 1152, 1280, 1536, 2304, 3456):
    w = (torch.randn(N, K, device="cuda") * 0.03).bfloat16(); b = torch.randn(N, device="cuda")
    out = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    f = lambda: ops.gemm(a, w, b, out, EPI_BIAS)
    for _ in range(5): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): f()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    c = lambda: torch.matmul(a, w.t(), out=out)
    for _ in range(3): c()
    e0.record()
    for _ in range(20): c()
    e1.record(); torch.cuda.synchronize()
    mc = e0.elapsed_time(e1) / 20
    print(f"K={K} N={N}: ours {ms:.4f} ms = {2.0*M*N*K/ms/1e9:.0f} TF/s | cuBLAS {mc:.4f} ms = {2.0*M*N*K/mc/1e9:.0f} TF/s | ratio {ms/mc:.3f}")
