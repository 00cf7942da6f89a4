#!/bin/bash
# ncu captures of the hot kernels (one block of FiT-XL/2 at the headline shape); plain run first, as the recipe requires.
mkdir -p gpurun_out
python tools/prof_kernels.py > gpurun_out/prof_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"${PROF_REGEX:-gemm_tc|attn_|ln_modulate}" -s ${PROF_SKIP:-9} -c ${PROF_COUNT:-9} -o gpurun_out/prof python tools/prof_kernels.py > gpurun_out/prof_ncu.log 2>&1
echo "ncu exit $?"; tail -5 gpurun_out/prof_ncu.log
