#!/bin/bash
# after a rewrite of cam_ctx_gate_kernel: GPU suite, per-kernel breakdown
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2g_pytest.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/r2g_pytest.log; tail -3 gpurun_out/r2g_pytest.log | cut -c1-300
timeout 300 python tools/kprof.py 250 bf16x3 > gpurun_out/r2g_kprof_bf16x3.txt 2>&1; grep -v "Warn\|_warn_once" gpurun_out/r2g_kprof_bf16x3.txt | head -14 | cut -c1-140
