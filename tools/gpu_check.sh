#!/bin/bash
# One gpurun call: GPU parity tests, a bench line, the per-kernel breakdown, and the ncu launch list.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
cat gpurun_out/bench.json
timeout 600 python bench.py --steps 5 --warmup 3 --precision bf16 --no-cpu-baseline > gpurun_out/bench_bf16.json 2> gpurun_out/bench_bf16.err
cat gpurun_out/bench_bf16.json
timeout 300 python tools/kprof.py 250 bf16x3 > gpurun_out/kprof_bf16x3.txt 2>&1
timeout 300 python tools/kprof.py 250 bf16 > gpurun_out/kprof_bf16.txt 2>&1
cat gpurun_out/kprof_bf16x3.txt
timeout 300 python tools/one_step.py 100 bf16x3 1 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv python tools/one_step.py 100 bf16x3 1 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
