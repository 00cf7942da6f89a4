#!/bin/bash
# One GPU-box visit: parity tests (each file in its own process so a CUDA fault in one does not mask the others), then a short bench.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
for f in test_gpu_kernels test_gpu_gemm test_gpu_model test_gpu_backward test_gpu_optim test_gpu_configs; do
  timeout 900 python -m pytest tests/$f.py -q -m gpu -x --timeout 600 > gpurun_out/$f.log 2>&1
  echo "$f exit $?" | tee -a gpurun_out/summary.txt
  tail -n 25 gpurun_out/$f.log
done
timeout 900 python bench.py --steps ${BENCH_STEPS:-10} --warmup 3 ${BENCH_ARGS} > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench exit $?" | tee -a gpurun_out/summary.txt
tail -c 3000 gpurun_out/bench.err; cat gpurun_out/bench.json
