#!/bin/bash
# short confirmations after the stem change: invariance tests / CLI tests (argument = test file stem)
mkdir -p gpurun_out
timeout ${2:-75} python -m pytest tests/test_gpu_${1:-invariance}.py -x -q -m gpu > gpurun_out/r2i_pytest_${1:-invariance}.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/r2i_pytest_${1:-invariance}.log; tail -3 gpurun_out/r2i_pytest_${1:-invariance}.log | cut -c1-300
