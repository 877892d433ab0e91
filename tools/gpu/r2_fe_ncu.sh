#!/bin/bash
mkdir -p gpurun_out
DEV=$PWD/fangyan_tts_b200/libfangyan_b200_dev.so
[ -f $DEV ] && export FY_LIB=$DEV
timeout 120 python tools/fe_one.py 1024 2 && timeout 600 ncu --set full --clock-control none --import-source on -k "regex:fbank2_kernel|logmel2_kernel|cmn_sub|logmel2_finish" -s 4 -c 4 -o gpurun_out/r2_fe2 python tools/fe_one.py 1024 2 > gpurun_out/r2_fe_ncu.log 2>&1
tail -2 gpurun_out/r2_fe_ncu.log
