#!/bin/bash
# gpurun call: S3 parity tests, kprof, then ncu --set full on selected kernels ($1 = regex, $2 = count)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_s3tok.py tests/test_gpu_golden.py -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/kprof.py 250 bf16x3 > gpurun_out/kprof_bf16x3.txt 2>&1; grep -A10 "kernel time total" gpurun_out/kprof_bf16x3.txt
timeout 300 python tools/one_step.py 100 bf16x3 1 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$1" -c ${2:-2} -o gpurun_out/prof_sel python tools/one_step.py 100 bf16x3 1 > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
