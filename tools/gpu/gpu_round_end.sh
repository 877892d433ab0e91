#!/bin/bash
# round-end evidence in one call: full GPU suite, smoke, reference arm, headline bench, ncu launch list of one step,
# ncu --set full of the attention kernel
bash tools/gpu/gpu_final.sh
bash tools/gpu/gpu_launchlist.sh
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:attention_tc" -s 2 -c 1 -o gpurun_out/attn_tc_final python tools/one_step.py 100 bf16x3 1 > gpurun_out/ncu_attn.log 2>&1; tail -1 gpurun_out/ncu_attn.log
