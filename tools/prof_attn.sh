mkdir -p gpurun_out
python tools/prof_kernels.py > gpurun_out/prof_plain_r1b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"attn_mma|ln_modulate" -s 3 -c 2 -f -o gpurun_out/prof_r1b python tools/prof_kernels.py > gpurun_out/prof_ncu_r1b.log 2>&1
echo "ncu exit $?"; tail -3 gpurun_out/prof_ncu_r1b.log
