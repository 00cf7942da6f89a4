#!/bin/bash
# Round-2 evidence trail (one GPU): (1) launch list of ONE sampling step (profiler-bracketed, no init kernels), (2) --set full capture of the
# block kernels of the build being benchmarked, (3) optional compute-sanitizer passes.  Each ncu command runs only after the same command
# exited 0 without ncu.
mkdir -p gpurun_out
TAG=${PROF_TAG:-r2a}
python tools/step_launches.py > gpurun_out/step_plain_$TAG.log 2>&1 &&
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$TAG.csv \
    python tools/step_launches.py > gpurun_out/step_ncu_$TAG.log 2>&1
echo "launch-list exit $?"; tail -1 gpurun_out/step_plain_$TAG.log
python tools/step_share.py gpurun_out/launches_$TAG.csv > gpurun_out/${TAG}_step_share.txt; head -14 gpurun_out/${TAG}_step_share.txt
python tools/prof_kernels.py > gpurun_out/prof_plain_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"${PROF_REGEX:-gemm_tc|attn_|ln_modulate|ln_apply}" -s ${PROF_SKIP:-9} -c ${PROF_COUNT:-9} -f -o gpurun_out/prof_$TAG \
    python tools/prof_kernels.py > gpurun_out/prof_ncu_$TAG.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/prof_ncu_$TAG.log
python tools/ncu_summary.py gpurun_out/prof_$TAG.ncu-rep gpurun_out/${TAG}_ncu_full_block_kernels.csv > /dev/null 2>&1; echo "summary exit $?"
if [ "${SANITIZE:-0}" = "1" ]; then
  for tool in racecheck synccheck; do
    timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_kernels.py -q -m gpu -x --timeout 1400 \
        -k "${SANITIZE_K:-not 4096}" > gpurun_out/${TAG}_sanitizer_$tool.log 2>&1
    echo "$tool exit $?"; tail -4 gpurun_out/${TAG}_sanitizer_$tool.log
  done
fi
