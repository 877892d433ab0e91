#!/bin/bash
# A/B of an env switch on the same box: $1 = VAR, runs tests with the default, then bench with VAR unset and VAR=0
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_s3tok.py tests/test_gpu_invariance.py tests/test_gpu_long.py tests/test_gpu_golden.py -q -x > gpurun_out/pytest_ab.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_ab.log
tail -4 gpurun_out/pytest_ab.log
bash tools/gpu/gpu_stalls.sh | grep -v "^exit"
for v in on 0 on 0; do
  if [ "$v" = "0" ]; then export $1=0; else unset $1; fi
  timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-bf16-mode > gpurun_out/bench_ab_$v.json 2> gpurun_out/bench_ab_$v.err; echo "$1=$v exit $?"
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_ab_$v.json'))
print(' ', round(d['ms_per_step'],2), round(d['value'],3), round(d['e2e']['value'],3), d['clocks']['sm_mhz'], round(d['roofline']['executed_frac'],4))
PY
done
