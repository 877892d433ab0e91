#!/bin/bash
# round 2, second call: attention role stalls (debug build), CAMPPlus / CLI / twin tests, twin latency, CLI end-to-end
mkdir -p gpurun_out
timeout 300 python tools/pp_stalls.py 250 bf16x3 2>&1 | grep -v "Warn\|_warn" | tee gpurun_out/r2_pp_stalls.txt
timeout 300 python tools/pp_stalls.py 64 bf16x3 30 2>&1 | grep -v "Warn\|_warn" | tee -a gpurun_out/r2_pp_stalls.txt
timeout 300 python tools/pp_stalls.py 250 bf16 2>&1 | grep -v "Warn\|_warn" | tee -a gpurun_out/r2_pp_stalls.txt
timeout 900 python -m pytest tests/test_gpu_campplus.py tests/test_gpu_cli.py tests/test_gpu_golden.py tests/test_gpu_invariance.py -x -q > gpurun_out/r2_second_pytest.log 2>&1; echo "pytest exit $?"
tail -8 gpurun_out/r2_second_pytest.log | cut -c1-600
timeout 300 python tools/twin_latency.py 5 > gpurun_out/r2_twin_latency.json 2> gpurun_out/r2_twin_latency.err; echo "twin exit $?"; tail -2 gpurun_out/r2_twin_latency.err | cut -c1-300; cut -c1-1200 gpurun_out/r2_twin_latency.json
timeout 300 python tools/kprof.py 250 bf16x3 > gpurun_out/r2_kprof.txt 2>&1; head -24 gpurun_out/r2_kprof.txt | cut -c1-150
timeout 900 python bench.py --cli-e2e 40000 > gpurun_out/r2_cli_e2e_b.json 2> gpurun_out/r2_cli_e2e_b.err; echo "cli exit $?"
tail -3 gpurun_out/r2_cli_e2e_b.err | cut -c1-400; cut -c1-3000 gpurun_out/r2_cli_e2e_b.json
