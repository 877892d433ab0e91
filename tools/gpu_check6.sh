#!/bin/bash
# gpurun call: try the CTA-pair GEMM under a short timeout first; fall back to FY_GEMM_2CTA=0 for the rest if it fails.
set -x
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_frontend.py -q -k gemm > gpurun_out/pytest_gemm.log 2>&1; rc=$?
echo "gemm test exit $rc" >> gpurun_out/pytest_gemm.log; tail -15 gpurun_out/pytest_gemm.log
if [ $rc -ne 0 ]; then export FY_GEMM_2CTA=0; echo "PAIR DISABLED"; nvidia-smi --query-gpu=name,memory.used --format=csv; fi
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
timeout 600 python tools/probe_precision.py > gpurun_out/probe_precision.txt 2>&1; tail -2 gpurun_out/probe_precision.txt
timeout 300 python tools/kprof.py 250 bf16x3 > gpurun_out/kprof_bf16x3.txt 2>&1; grep -A12 "kernel time total" gpurun_out/kprof_bf16x3.txt
FY_GEMM_2CTA=0 timeout 300 python tools/kprof.py 250 bf16x3 > gpurun_out/kprof_bf16x3_1cta.txt 2>&1; grep -A3 "kernel time total" gpurun_out/kprof_bf16x3_1cta.txt
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json
