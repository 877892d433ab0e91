#!/bin/bash
# gpurun call: GPU tests, role-stall profile (debug build), kprof, bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
bash tools/gpu/gpu_stalls.sh
timeout 300 python tools/kprof.py 250 bf16x3 > gpurun_out/kprof_bf16x3.txt 2>&1; tail -26 gpurun_out/kprof_bf16x3.txt
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json
