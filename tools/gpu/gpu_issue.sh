#!/bin/bash
# converged-warp MMA issue (GEMM + tcgen05 attention): full GPU suite at the default, attention-TC subset, kernel
# breakdowns for both attention kernels, role stalls, A/B bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log | cut -c1-250
FY_ATTN_TC=1 timeout 600 python -m pytest tests/test_gpu_s3tok.py tests/test_gpu_invariance.py tests/test_gpu_long.py tests/test_gpu_golden.py -q -x > gpurun_out/pytest_attn_tc.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_attn_tc.log
tail -4 gpurun_out/pytest_attn_tc.log | cut -c1-250
FY_ATTN_TC=0 timeout 300 python tools/kprof.py 2>&1 | grep -v "Warn\|_warn" | head -12 | tee gpurun_out/kprof_attn_0.txt
FY_ATTN_TC=1 timeout 300 python tools/kprof.py 2>&1 | grep -v "Warn\|_warn" | head -12 | tee gpurun_out/kprof_attn_1.txt
[ -f fangyan_tts_b200/libfangyan_b200_prof.so ] && timeout 300 python tools/attn_stalls.py 250 bf16x3 2>&1 | grep -v "Warn\|_warn" | tee gpurun_out/attn_stalls.txt
for v in 0 1 0 1; do
  FY_ATTN_TC=$v timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-bf16-mode > gpurun_out/bench_attn_$v.json 2> gpurun_out/bench_attn_$v.err; echo "FY_ATTN_TC=$v exit $?"
  python - <<PY
import json
d=json.load(open('gpurun_out/bench_attn_$v.json'))
print(' ', round(d['ms_per_step'],2), round(d['value'],3), round(d['e2e']['value'],3), d['clocks']['sm_mhz'], round(d['roofline']['executed_frac'],4))
PY
done
