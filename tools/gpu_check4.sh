#!/bin/bash
# gpurun call: GPU tests, precision probe, kprof, bench, ncu launch list (gpu__time_duration only).
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
timeout 600 python tests/manual/probe_precision.py > gpurun_out/probe_precision.txt 2>&1; tail -3 gpurun_out/probe_precision.txt
timeout 300 python tools/kprof.py 250 bf16x3 > gpurun_out/kprof_bf16x3.txt 2>&1; tail -26 gpurun_out/kprof_bf16x3.txt
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json
timeout 300 python tools/one_step.py 100 bf16x3 1 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv python tools/one_step.py 100 bf16x3 1 > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
