#!/bin/bash
# FY_ATTN_TC=1 (tcgen05 attention): parity under a short timeout first, then kernel breakdown
mkdir -p gpurun_out
export FY_ATTN_TC=1
timeout 200 python -m pytest tests/test_gpu_s3tok.py tests/test_gpu_invariance.py -q -x > gpurun_out/pytest_tc.log 2>&1; rc=$?; echo "pytest exit $rc" >> gpurun_out/pytest_tc.log
tail -6 gpurun_out/pytest_tc.log | cut -c1-200
if [ $rc -ne 0 ]; then exit 0; fi
timeout 300 python tools/kprof.py 250 bf16x3 2>&1 | grep -v "Warn\|_warn" | head -5
