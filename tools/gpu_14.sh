#!/bin/bash
# ncu --set full of the three kernels that carry the step: tcgen05 GEMM (roofline shape), wgrad, BN apply
mkdir -p gpurun_out
for w in fprop wgrad bn_apply; do
  timeout 120 python tools/one_kernel.py --which $w > gpurun_out/one_$w.log 2>&1 || { echo "plain $w failed"; cat gpurun_out/one_$w.log | tail -3; continue; }
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_tc_gemm -s 2 -c 2 -f -o gpurun_out/prof_gemm python tools/one_kernel.py --which fprop > gpurun_out/ncu_gemm.log 2>&1; echo "ncu gemm exit=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_tc_wgrad -s 2 -c 2 -f -o gpurun_out/prof_wgrad python tools/one_kernel.py --which wgrad > gpurun_out/ncu_wgrad.log 2>&1; echo "ncu wgrad exit=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:bn_apply -s 2 -c 2 -f -o gpurun_out/prof_bn_apply python tools/one_kernel.py --which bn_apply > gpurun_out/ncu_bn.log 2>&1; echo "ncu bn exit=$?"
ls -la gpurun_out/*.ncu-rep
