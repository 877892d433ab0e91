#!/bin/bash
# gpurun call: ncu --set full on selected kernels of one step ($1 regex, $2 skip, $3 count, $4 output name)
mkdir -p gpurun_out
timeout 300 python tools/one_step.py 100 bf16x3 1 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$1" -s ${2:-0} -c ${3:-4} -o gpurun_out/${4:-prof_sel} python tools/one_step.py 100 bf16x3 1 > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log
