"""ncu target: the LN-prologue SwiGLU GEMM alone at the headline shape (M = 131072 rows, K = 1152, N = 6144).
gemm_tc_kernel launches: #0 the c-vector GEMM, #1.. the fused GEMM."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fit_b200 import ops
from fit_b200._lib import EPI_BIAS, EPI_SWIGLU

M, K, Hp, T = int(os.environ.get("ROWS", 131072)), 1152, 3072, 256
S = M // T
g = torch.Generator(device="cuda").manual_seed(0)
x = torch.randn(M, K, device="cuda", generator=g).bfloat16()
W = (torch.randn(2 * Hp, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
b = torch.randn(2 * Hp, device="cuda", generator=g) * 0.1
mod = torch.randn(S, 3 * K, device="cuda", generator=g) * 0.3
packed = torch.empty(1, 2 * S, K, dtype=torch.bfloat16, device="cuda")
ops.modulation_pack(mod, mod[:, K:], 3 * K, S, K, 1, 3 * K, packed)
c = torch.empty(2 * S, 2 * Hp, dtype=torch.float32, device="cuda")
ops.gemm(packed[0], W, b, c, EPI_BIAS, bias_row0=S)
parts = torch.empty(M, 10, 2, dtype=torch.float32, device="cuda")
ops.row_stats(x, parts)
out = torch.empty(M, Hp, dtype=torch.bfloat16, device="cuda")
for _ in range(int(os.environ.get("REPS", 3))):
    ops.gemm(x, W, None, out, EPI_SWIGLU, rows_per_sample=T, ln=(packed[0, :S], c, parts, 1e-6))
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
ops.gemm(x, W, None, out, EPI_SWIGLU, rows_per_sample=T, ln=(packed[0, :S], c, parts, 1e-6))
e1.record(); torch.cuda.synchronize()
print("fused swiglu gemm %.3f ms" % e0.elapsed_time(e1))
