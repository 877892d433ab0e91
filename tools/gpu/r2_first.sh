#!/bin/bash
# round 2, first call: GPU suite, headline bench with parity, CLI end-to-end, long-utterance stress
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader | head -2
df -h /dev/shm | tail -1
nproc
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/r2_pytest.log
tail -5 gpurun_out/r2_pytest.log | cut -c1-300
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; echo "bench exit $?"
tail -3 gpurun_out/r2_bench.err; cut -c1-1500 gpurun_out/r2_bench.json
timeout 900 python bench.py --cli-e2e 20000 > gpurun_out/r2_cli_e2e.json 2> gpurun_out/r2_cli_e2e.err; echo "cli exit $?"
tail -5 gpurun_out/r2_cli_e2e.err | cut -c1-400; cut -c1-3000 gpurun_out/r2_cli_e2e.json
timeout 600 python bench.py --long-stress --steps 3 > gpurun_out/r2_long.json 2> gpurun_out/r2_long.err; echo "long exit $?"
tail -3 gpurun_out/r2_long.err | cut -c1-400; cut -c1-2500 gpurun_out/r2_long.json
