#!/bin/bash
# attention A/B: the side-by-side build (FY_LIB=..._dev.so when present, else the default build) against the kernels it replaces
mkdir -p gpurun_out
DEV=$PWD/fangyan_tts_b200/libfangyan_b200_dev.so
[ -f $DEV ] && export FY_LIB=$DEV
echo "lib: ${FY_LIB:-default}"
timeout 900 python -m pytest tests/test_gpu_s3tok.py tests/test_gpu_invariance.py -x -q > gpurun_out/r2_attn_pytest.log 2>&1; echo "pytest exit $?"
tail -6 gpurun_out/r2_attn_pytest.log | cut -c1-400
for v in 2 1; do
  FY_ATTN_TC=$v timeout 300 python tools/kprof.py 250 bf16x3 > gpurun_out/r2_kprof_attn_$v.txt 2>&1
  echo "FY_ATTN_TC=$v"; grep -E "attention|kernel time" gpurun_out/r2_kprof_attn_$v.txt | cut -c1-160
done
for v in 2 0; do
  FY_ATTN_TC=$v timeout 300 python tools/kprof.py 250 bf16 > gpurun_out/r2_kprof_attn_bf16_$v.txt 2>&1
  echo "bf16 FY_ATTN_TC=$v"; grep -E "attention|kernel time" gpurun_out/r2_kprof_attn_bf16_$v.txt | cut -c1-160
done
timeout 600 python -m pytest tests/test_gpu_parity_config1.py -x -q > gpurun_out/r2_attn_parity.log 2>&1; echo "parity exit $?"; tail -3 gpurun_out/r2_attn_parity.log | cut -c1-300
