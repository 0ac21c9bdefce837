#!/bin/bash
mkdir -p gpurun_out
SH="256,56,64,256,1"
timeout 120 python tools/one_kernel.py --which fprop --shape $SH > gpurun_out/one_mb.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_tc_gemm -s 2 -c 1 -f -o gpurun_out/prof_gemm_64_256 python tools/one_kernel.py --which fprop --shape $SH > gpurun_out/ncu_mb.log 2>&1; echo "ncu exit=$?"
SH2="256,56,64,64,3"
timeout 120 python tools/one_kernel.py --which fprop --shape $SH2 > gpurun_out/one_mb2.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_tc_gemm -s 2 -c 1 -f -o gpurun_out/prof_gemm_64_64_3x3 python tools/one_kernel.py --which fprop --shape $SH2 > gpurun_out/ncu_mb2.log 2>&1; echo "ncu exit=$?"
