#!/bin/bash
# which part of the tcgen05 attention makes the full-size repeatability test fail? default / never-rescale / always-rescale
mkdir -p gpurun_out
export FY_ATTN_TC=1
for tag in "" _noslack _slack0; do
  FY_LIB=$PWD/fangyan_tts_b200/libfangyan_b200$tag.so timeout 300 python -m pytest tests/test_gpu_long.py -q -x -k full_size 2>&1 | tail -3 | cut -c1-200
done
bash tools/gpu/gpu_ncu.sh attention_tc 2 1 attn_v3
