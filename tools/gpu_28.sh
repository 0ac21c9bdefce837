#!/bin/bash
mkdir -p gpurun_out
J="dw_fprop:256,56,128,1;dw_fprop:256,14,512,1"
timeout 120 python tools/one_kernel.py --jobs "$J" > gpurun_out/ok.log 2>&1 || { echo plain run failed; tail -5 gpurun_out/ok.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"dw_fwd" -o gpurun_out/prof_dw2 -f \
   python tools/one_kernel.py --iters 1 --jobs "$J" > gpurun_out/ncu.log 2>&1
echo "ncu exit=$?"; tail -3 gpurun_out/ncu.log
