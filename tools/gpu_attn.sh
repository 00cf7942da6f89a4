#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -q -m gpu -x -k attention > gpurun_out/attn_test.log 2>&1; echo "attn test exit $?"; tail -12 gpurun_out/attn_test.log
timeout 300 python -m pytest tests/test_gpu_gemm.py -q -m gpu -x -k rope > gpurun_out/rope_test.log 2>&1; echo "rope test exit $?"; tail -12 gpurun_out/rope_test.log
timeout 120 python tools/bench_attn.py; VALID=128 timeout 120 python tools/bench_attn.py; HD=64 timeout 120 python tools/bench_attn.py; PRE=0 timeout 120 python tools/bench_attn.py
