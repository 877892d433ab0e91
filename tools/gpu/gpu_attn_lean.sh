#!/bin/bash
# short check of an attention-kernel change (no bench): parity subset, kernel breakdown, role stalls.
# $1 = optional build tag of a side-by-side library (FY_BUILD_TAG=$1 and ${1}prof), loaded through FY_LIB
mkdir -p gpurun_out
if [ -n "$1" ]; then export FY_LIB=$PWD/fangyan_tts_b200/libfangyan_b200_$1.so; PROF=$PWD/fangyan_tts_b200/libfangyan_b200_${1}prof.so; else PROF=$PWD/fangyan_tts_b200/libfangyan_b200_prof.so; fi
timeout 150 python -m pytest tests/test_gpu_s3tok.py -q -x > gpurun_out/pytest_attn_first.log 2>&1 || { echo "first parity run failed"; tail -30 gpurun_out/pytest_attn_first.log | cut -c1-250; exit 1; }
timeout 400 python -m pytest tests/test_gpu_invariance.py tests/test_gpu_long.py tests/test_gpu_golden.py -q -x > gpurun_out/pytest_attn_tc.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_attn_tc.log
tail -6 gpurun_out/pytest_attn_tc.log | cut -c1-250
timeout 300 python tools/kprof.py 2>&1 | grep -v "Warn\|_warn" | head -8 | tee gpurun_out/kprof_attn_tc.txt
[ -f $PROF ] && FY_LIB=$PROF timeout 300 python tools/attn_stalls.py 250 bf16x3 2>&1 | grep -v "Warn\|_warn" | tee gpurun_out/attn_stalls.txt
