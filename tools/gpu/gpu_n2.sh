#!/bin/bash
# gpurun --gpus 2: attention-kernel agreement test (incl. the forced-rescale variant), multi-GPU CLI test, torchrun bench at 2 ranks
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_s3tok.py tests/test_gpu_cli.py -q -x > gpurun_out/pytest_n2.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_n2.log
tail -4 gpurun_out/pytest_n2.log | cut -c1-250
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu-baseline --no-bf16-mode > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench exit $?"
tail -2 gpurun_out/bench_n2.err; cut -c1-400 gpurun_out/bench_n2.json
