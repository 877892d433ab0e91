#!/bin/bash
# CAMPPlus: BatchNorm+ReLU folded into the bottleneck / transit GEMMs (side-by-side build) against the separate pass; ncu
# --set full of the four layer GEMMs at the bench batch size (DRAM traffic for the roofline object)
mkdir -p gpurun_out
DEV=$PWD/fangyan_tts_b200/libfangyan_b200_dev.so
timeout 600 ncu --set full --clock-control none -k "regex:gemm_kernel" -s 2 -c 4 -o gpurun_out/r2_gemm_pair_bench python tools/s2_one.py 370 bf16x3 > gpurun_out/r2_ncu_gemm.log 2>&1; tail -2 gpurun_out/r2_ncu_gemm.log
[ -f $DEV ] && export FY_LIB=$DEV
echo "lib: ${FY_LIB:-default}"
timeout 900 python -m pytest tests/test_gpu_campplus.py tests/test_gpu_golden.py tests/test_gpu_invariance.py tests/test_gpu_cli.py tests/test_gpu_long.py -x -q 2>&1 | tail -6 | cut -c1-400
for v in 1 0; do
  FY_CAM_FUSED_BN=$v timeout 300 python tools/kprof.py 250 bf16x3 > gpurun_out/r2_kprof_cam_$v.txt 2>&1
  echo "FY_CAM_FUSED_BN=$v"; grep -E "kernel time|bnrelu|gemm_kernel<128|gemm_kernel<256, 2|gemm_kernel<32" gpurun_out/r2_kprof_cam_$v.txt | cut -c1-140
done
