#!/bin/bash
# tcgen05 attention vs mma.sync attention on the same box: parity suite with FY_ATTN_TC=1, then alternating bench runs
mkdir -p gpurun_out
FY_ATTN_TC=1 timeout 900 python -m pytest tests/test_gpu_s3tok.py tests/test_gpu_invariance.py tests/test_gpu_long.py tests/test_gpu_golden.py -q -x > gpurun_out/pytest_attn_tc.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_attn_tc.log
tail -4 gpurun_out/pytest_attn_tc.log
for v in 0 1 0 1; do
  FY_ATTN_TC=$v timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-bf16-mode > gpurun_out/bench_attn_$v.json 2> gpurun_out/bench_attn_$v.err; echo "FY_ATTN_TC=$v exit $?"
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_attn_$v.json'))
print(' ', round(d['ms_per_step'],2), round(d['value'],3), round(d['e2e']['value'],3), d['clocks']['sm_mhz'], round(d['roofline']['executed_frac'],4))
PY
done
FY_ATTN_TC=1 timeout 300 python tools/kprof.py 2>&1 | grep -v "Warn\|_warn" | tail -40 > gpurun_out/kprof_attn_tc.txt; tail -30 gpurun_out/kprof_attn_tc.txt
