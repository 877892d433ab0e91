#!/bin/bash
# gpurun --gpus N: long-utterance test, front-end sweep (rank 0 only), torchrun bench at N ranks
N=${1:-4}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_long.py tests/test_gpu_invariance.py tests/test_gpu_cli.py -q > gpurun_out/pytest_multi.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_multi.log
tail -6 gpurun_out/pytest_multi.log
timeout 300 python tools/bench_frontend.py > gpurun_out/frontend_sweep.jsonl 2> gpurun_out/frontend_sweep.err; cat gpurun_out/frontend_sweep.jsonl | cut -c1-420
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench exit $?"
tail -2 gpurun_out/bench_n$N.err; cat gpurun_out/bench_n$N.json | cut -c1-300
