#!/bin/bash
# gpurun --gpus N: drop-in CLIs end to end over 200k keys on N GPUs (lazy NCCL communicator, array-direct .pt writers)
N=${1:-8}; K=${2:-200000}
mkdir -p gpurun_out
timeout 420 python bench.py --gpus $N --cli-e2e $K > gpurun_out/r2_cli_e2e_n$N.json 2> gpurun_out/r2_cli_e2e_n$N.err; echo "cli exit $?"
tail -5 gpurun_out/r2_cli_e2e_n$N.err | cut -c1-400; cut -c1-3500 gpurun_out/r2_cli_e2e_n$N.json
