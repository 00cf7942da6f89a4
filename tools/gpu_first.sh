#!/bin/bash
# First GPU visit of a session: smoke, parity tests, bench, ncu launch list and --set full capture.
mkdir -p gpurun_out
TAG=${PROF_TAG:-r1}
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$TAG.log 2>&1; echo "smoke exit $?" | tee gpurun_out/summary.txt
bash tools/gpu_round.sh
bash tools/gpu_prof.sh
