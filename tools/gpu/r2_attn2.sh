#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_s3tok.py tests/test_gpu_invariance.py tests/test_gpu_long.py -x -q > gpurun_out/r2_attn2_pytest.log 2>&1; echo "pytest exit $?"
tail -4 gpurun_out/r2_attn2_pytest.log | cut -c1-400
timeout 300 python tools/pp_stalls.py 250 bf16x3 2>&1 | grep -v "Warn\|_warn" | tee gpurun_out/r2_pp_stalls2.txt
timeout 300 python tools/pp_stalls.py 64 bf16x3 30 2>&1 | grep -v "Warn\|_warn" | tee -a gpurun_out/r2_pp_stalls2.txt
for prec in bf16x3 bf16; do
  timeout 300 python tools/kprof.py 250 $prec > gpurun_out/r2_kprof2_$prec.txt 2>&1
  grep -E "attention|kernel time" gpurun_out/r2_kprof2_$prec.txt | cut -c1-160
done
