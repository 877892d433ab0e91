#!/bin/bash
# gpurun --gpus 2: multi-GPU CLI test (NCCL gather to rank 0), torchrun bench at 2 ranks with the exchange inside the step
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_cli.py -q -x > gpurun_out/r2_pytest_n2.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2_pytest_n2.log
tail -4 gpurun_out/r2_pytest_n2.log | cut -c1-300
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo "bench exit $?"
tail -3 gpurun_out/r2_bench_n2.err | cut -c1-300; cut -c1-1800 gpurun_out/r2_bench_n2.json
python -c "
import json; d=json.load(open('gpurun_out/r2_bench_n2.json')); print(json.dumps({k: d.get(k) for k in ('value','ms_per_step','gather','parity','e2e')})[:1500])"
timeout 600 python bench.py --gpus 2 --cli-e2e 40000 > gpurun_out/r2_cli_e2e_n2.json 2> gpurun_out/r2_cli_e2e_n2.err; echo "cli exit $?"; tail -3 gpurun_out/r2_cli_e2e_n2.err | cut -c1-300; cut -c1-2500 gpurun_out/r2_cli_e2e_n2.json
