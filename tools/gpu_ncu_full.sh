#!/bin/bash
mkdir -p gpurun_out
J="fprop:256,14,256,256,3;wgrad_acc:256,56,64,256,1;wgrad_acc:256,14,256,256,3;wgrad_acc:256,56,64,64,3"
timeout 120 python tools/one_kernel.py --jobs "$J" > gpurun_out/ok.log 2>&1 || { echo plain run failed; tail -5 gpurun_out/ok.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"conv_tc" -o gpurun_out/prof_r01_v9 -f \
   python tools/one_kernel.py --iters 2 --jobs "$J" > gpurun_out/ncu.log 2>&1
echo "ncu exit=$?"; tail -3 gpurun_out/ncu.log
