#!/bin/bash
# quick loop: S3 + invariance tests, kernel breakdown at 250 utterances
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_s3tok.py tests/test_gpu_invariance.py tests/test_gpu_campplus.py -q -x > gpurun_out/pytest_quick.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_quick.log
tail -4 gpurun_out/pytest_quick.log
timeout 300 python tools/kprof.py 250 bf16x3 > gpurun_out/kprof_bf16x3.txt 2>&1; grep -v "Warn\|_warn" gpurun_out/kprof_bf16x3.txt | head -14
