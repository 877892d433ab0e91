#!/bin/bash
# ncu launch list (gpu__time_duration only) of one device-resident step at 100 utterances, after a plain run exits 0
mkdir -p gpurun_out
timeout 300 python tools/one_step.py 100 bf16x3 1 > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches.csv python tools/one_step.py 100 bf16x3 1 > gpurun_out/ncu.log 2>&1
tail -2 gpurun_out/ncu.log; wc -l gpurun_out/launches.csv
