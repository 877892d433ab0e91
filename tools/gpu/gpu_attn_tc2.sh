#!/bin/bash
mkdir -p gpurun_out
export FY_ATTN_TC=1
timeout 600 python -m pytest tests/test_gpu_invariance.py tests/test_gpu_long.py tests/test_gpu_golden.py tests/test_gpu_s3tok.py tests/test_gpu_cli.py -q 2>&1 | tail -12 | cut -c1-220
