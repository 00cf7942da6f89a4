#!/bin/bash
# training-path visit: optimizer / backward / config tests, then the bench line with the extras (train step, strong scaling, eager torch)
mkdir -p gpurun_out
for f in test_gpu_optim test_gpu_backward test_gpu_configs; do
  timeout 600 python -m pytest tests/$f.py -q -m gpu -x --timeout 300 > gpurun_out/tr_$f.log 2>&1
  echo "$f exit $?"; tail -n 5 gpurun_out/tr_$f.log
done
timeout 300 python -m pytest tests/test_gpu_model.py -q -m gpu -x --timeout 200 -k "ln_prologue or tiny_forward" > gpurun_out/tr_model.log 2>&1; echo "model exit $?"; tail -3 gpurun_out/tr_model.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/tr_bench.json 2> gpurun_out/tr_bench.err; echo "bench exit $?"
python - <<P
import json
d = json.loads(open("gpurun_out/tr_bench.json").read().strip().splitlines()[-1])
print({k: d[k] for k in ("value", "ms_per_step")}, d["clocks"])
print(json.dumps(d["extra"], indent=0)[:1500])
P
