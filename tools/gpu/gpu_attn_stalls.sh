#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/attn_stalls.py 250 bf16x3 2>&1 | grep -v "Warn\|_warn" | tee gpurun_out/attn_stalls.txt
