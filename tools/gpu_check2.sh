#!/bin/bash
# gpurun call: full GPU test suite, precision probe, bench, then ncu --set full on the dominant GEMM.
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python tools/probe_precision.py > gpurun_out/probe_precision.txt 2>&1; cat gpurun_out/probe_precision.txt | tail -4
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
timeout 300 python tools/one_step.py 100 bf16x3 1 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 121 -c 5 -o gpurun_out/prof_gemm python tools/one_step.py 100 bf16x3 1 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
