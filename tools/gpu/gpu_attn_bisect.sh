#!/bin/bash
# does the full-size repeatability test depend on the in-TMEM rescale path of the tcgen05 attention?
# default slack (2^8) / never rescale / rescale whenever a tile raises the row maximum
mkdir -p gpurun_out
for slack in 8 1e30 0; do
  echo "FY_ATTN_SLACK=$slack"
  FY_ATTN_SLACK=$slack timeout 300 python -m pytest tests/test_gpu_long.py -q -x -k full_size 2>&1 | tail -3 | cut -c1-200
done
