"""Time the four block GEMMs of FiT-XL/2 at the headline shape (M = 512 x 256 rows), back to back (sustained clocks) and alone with an
L2 flush in between.  FITB200_GEMM_PAIR=0/1 selects single-CTA / CTA-pair tiles (read once per process)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fit_b200 import ops
from fit_b200._lib import EPI_BIAS, EPI_BIAS_GATE_RESID, EPI_BIAS_ROPE, EPI_SWIGLU

M, D, H, T, R = int(os.environ.get("ROWS", 512)) * 256, 1152, 3072, 256, int(os.environ.get("ROWS", 512))
dev = "cuda"
bf = torch.bfloat16
g = torch.Generator(device=dev).manual_seed(0)
rn = lambda *s: torch.randn(*s, device=dev, generator=g)
h = rn(M, D).to(bf); att = rn(M, D).to(bf); gg = rn(M, H).to(bf); xs = rn(M, D).to(bf)
qkv = torch.empty(M, 3 * D, dtype=bf, device=dev); gout = torch.empty(M, H, dtype=bf, device=dev)
w_qkv = (rn(3 * D, D) * 0.03).to(bf); b_qkv = rn(3 * D) * 0.02
w_proj = (rn(D, D) * 0.03).to(bf); b_proj = rn(D) * 0.02
w_fc1 = (rn(2 * H, D) * 0.03).to(bf); b_fc1 = rn(2 * H) * 0.02
w_fc2 = (rn(D, H) * 0.02).to(bf); b_fc2 = rn(D) * 0.02
gate = rn(R, D) * 0.1
pos = rn(1, T, 72)
cases = {
    "swiglu": (lambda: ops.gemm(h, w_fc1, b_fc1, gout, EPI_SWIGLU), 2.0 * M * D * 2 * H),
    "qkv+rope": (lambda: ops.gemm(h, w_qkv, b_qkv, qkv, EPI_BIAS_ROPE, rope=pos, rope_cols=2 * D), 2.0 * M * D * 3 * D),
    "qkv": (lambda: ops.gemm(h, w_qkv, b_qkv, qkv, EPI_BIAS), 2.0 * M * D * 3 * D),
    "fc2+gate": (lambda: ops.gemm(gg, w_fc2, b_fc2, xs, EPI_BIAS_GATE_RESID, gate=gate, ld_gate=D, rows_per_sample=T, resid=xs), 2.0 * M * H * D),
    "proj+gate": (lambda: ops.gemm(att, w_proj, b_proj, xs, EPI_BIAS_GATE_RESID, gate=gate, ld_gate=D, rows_per_sample=T, resid=xs), 2.0 * M * D * D),
}
scratch = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device=dev)
# box normaliser: the cuBLAS bf16 GEMM of the same shape (no epilogue), interleaved with ours in rounds (power / thermal state drifts)
ref = {"swiglu": (h, w_fc1), "qkv+rope": (h, w_qkv), "qkv": (h, w_qkv), "fc2+gate": (gg, w_fc2), "proj+gate": (att, w_proj)}


def timed(fn, n=20):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


print("pair =", os.environ.get("FITB200_GEMM_PAIR", "1"))
for name, (fn, fl) in cases.items():
    A, Wt = ref[name]
    o = torch.empty(A.shape[0], Wt.shape[0], dtype=bf, device=dev)
    cb = lambda: torch.matmul(A, Wt.t(), out=o)
    fn(); cb(); torch.cuda.synchronize()
    ours, cub = [], []
    for _ in range(7):
        cub.append(timed(cb)); ours.append(timed(fn))
    del o
    ratios = sorted(a / b for a, b in zip(ours, cub))
    ours.sort(); cub.sort()
    tb = ours[len(ours) // 2]
    print(f"{name:10s} {tb:.4f} ms = {fl / tb / 1e9:7.1f} TF/s | cuBLAS plain GEMM {cub[len(cub) // 2]:.4f} ms | ours/cuBLAS time: median {ratios[len(ratios) // 2]:.3f} (min {ratios[0]:.3f}, max {ratios[-1]:.3f})")
