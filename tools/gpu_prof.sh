#!/bin/bash
# ncu evidence, per the B200 profiling recipe: (1) launch list of a short bench run, (2) --set full capture of the hot kernels
# of one FiT-XL/2 block at the headline shape.  Each ncu command runs only after the same command exited 0 without ncu.
mkdir -p gpurun_out
TAG=${PROF_TAG:-r1}
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_plain_$TAG.json 2> gpurun_out/bench_plain_$TAG.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -s ${LIST_SKIP:-220} -c ${LIST_COUNT:-420} --csv --log-file gpurun_out/launches_$TAG.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_ncu_$TAG.log 2>&1
echo "launch-list exit $?"
python tools/prof_kernels.py > gpurun_out/prof_plain_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"${PROF_REGEX:-gemm_tc|attn_|ln_modulate}" -s ${PROF_SKIP:-9} -c ${PROF_COUNT:-9} -f -o gpurun_out/prof_$TAG \
    python tools/prof_kernels.py > gpurun_out/prof_ncu_$TAG.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/prof_ncu_$TAG.log
