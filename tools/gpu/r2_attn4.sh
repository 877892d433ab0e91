#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_s3tok.py tests/test_gpu_invariance.py tests/test_gpu_parity_config1.py -x -q -s 2>&1 | grep -E "n_layer=|passed|failed|token_match|Error|assert" | cut -c1-400 | tee gpurun_out/r2_attn4_pytest.log
timeout 300 python tools/pp_stalls.py 250 bf16x3 2>&1 | grep -v "Warn\|_warn" | tee gpurun_out/r2_pp_stalls4.txt
timeout 300 python tools/pp_stalls.py 64 bf16x3 30 2>&1 | grep -v "Warn\|_warn" | tee -a gpurun_out/r2_pp_stalls4.txt
timeout 300 python tools/kprof.py 250 bf16x3 > gpurun_out/r2_kprof4.txt 2>&1
grep -E "attention|kernel time" gpurun_out/r2_kprof4.txt | cut -c1-160
