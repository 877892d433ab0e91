#!/bin/bash
# after the 8-rows-per-block stem kernel: CAMPPlus + golden + config-1 parity tests, per-kernel breakdown
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_campplus.py tests/test_gpu_golden.py tests/test_gpu_parity_config1.py -x -q -m gpu > gpurun_out/r2s_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/r2s_pytest.log; tail -3 gpurun_out/r2s_pytest.log | cut -c1-300
timeout 100 python tools/kprof.py 250 bf16x3 > gpurun_out/r2s_kprof_bf16x3.txt 2>&1; grep -v "Warn\|_warn_once" gpurun_out/r2s_kprof_bf16x3.txt | head -12 | cut -c1-140
