#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/pp_stalls.py 250 bf16x3 2>&1 | grep -v "Warn\|_warn" | tee gpurun_out/r2_pp_stalls.txt
timeout 300 python tools/pp_stalls.py 64 bf16x3 30 2>&1 | grep -v "Warn\|_warn" | tee -a gpurun_out/r2_pp_stalls.txt
timeout 300 python tools/pp_stalls.py 250 bf16 2>&1 | grep -v "Warn\|_warn" | tee -a gpurun_out/r2_pp_stalls.txt
timeout 600 python -m pytest tests/test_gpu_s3tok.py -x -q -k "epsilon or matches" 2>&1 | tail -3
