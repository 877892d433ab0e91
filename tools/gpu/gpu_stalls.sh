#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/gemm_stalls.py 250 bf16x3 > gpurun_out/gemm_stalls.txt 2>&1; echo "exit $?" >> gpurun_out/gemm_stalls.txt
timeout 300 python tools/gemm_stalls.py 250 bf16 >> gpurun_out/gemm_stalls.txt 2>&1; echo "exit $?" >> gpurun_out/gemm_stalls.txt
cat gpurun_out/gemm_stalls.txt
