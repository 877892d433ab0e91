#!/bin/bash
# fy_fbank_cmn_static: its test alone in a fresh process, then the whole GPU suite (the test runs last), then smoke
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -x -m gpu > gpurun_out/r2_noalloc_alone.log 2>&1; echo "alone exit $?"; tail -5 gpurun_out/r2_noalloc_alone.log | cut -c1-300
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/r2_pytest_gpu.log 2>&1; echo "suite exit $?"; tail -6 gpurun_out/r2_pytest_gpu.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
