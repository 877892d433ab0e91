#!/bin/bash
# gpurun --gpus 8: torchrun bench at 8 ranks (phase markers + watchdog on stderr), nothing else -- 8x the box time
N=${1:-8}
mkdir -p gpurun_out
FY_BENCH_WATCHDOG=${FY_BENCH_WATCHDOG:-120} timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err; echo "bench exit $?"
grep -c "" gpurun_out/r2_bench_n$N.err; grep "bench r0\|bench r7" gpurun_out/r2_bench_n$N.err | cut -c1-200 | tail -40
grep -v "^\[bench" gpurun_out/r2_bench_n$N.err | tail -60 | cut -c1-240
grep "^{" gpurun_out/r2_bench_n$N.json | tail -1 | cut -c1-1500
