#!/bin/bash
# LN-prologue visit: the new GEMM tests, the model-level parity tests through the fused path, then fused vs un-fused bench lines.
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_gpu_gemm.py -q -m gpu -x --timeout 120 -k "ln_prologue or row_sums" > gpurun_out/ln_gemm.log 2>&1
echo "ln gemm tests exit $?"; tail -n 30 gpurun_out/ln_gemm.log
for f in ${LN_TESTS-test_gpu_gemm test_gpu_model test_gpu_configs test_gpu_kernels}; do
  timeout 900 python -m pytest tests/$f.py -q -m gpu --timeout 600 > gpurun_out/ln_$f.log 2>&1
  echo "$f exit $?"; tail -n 6 gpurun_out/ln_$f.log
done
for fuse in ${LN_FUSE-1 0}; do
  FITB200_LN_FUSE=$fuse timeout 600 python bench.py --steps 10 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/ln_bench_fuse$fuse.json 2> gpurun_out/ln_bench_fuse$fuse.err
  echo "bench fuse=$fuse exit $?"; python - <<P
import json
try:
    d = json.loads(open("gpurun_out/ln_bench_fuse$fuse.json").read().strip().splitlines()[-1])
    print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches")}, d["clocks"])
    for k, v in d["roofline"].get("kernels", {}).items():
        print("  ", k, {kk: (round(vv, 4) if isinstance(vv, float) else vv) for kk, vv in v.items() if kk in ("ms", "ms_mean", "frac")})
except Exception as e:
    print("no bench line", e); print(open("gpurun_out/ln_bench_fuse$fuse.err").read()[-2000:])
P
done
