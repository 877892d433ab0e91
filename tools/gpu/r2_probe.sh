#!/bin/bash
mkdir -p gpurun_out
timeout 400 python tools/micro/startup_probe.py ${1:-8} > gpurun_out/r2_startup_probe.jsonl 2> gpurun_out/r2_startup_probe.err; echo "exit $?"
cat gpurun_out/r2_startup_probe.jsonl; tail -5 gpurun_out/r2_startup_probe.err | cut -c1-300
