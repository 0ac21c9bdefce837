#!/bin/bash
# Round-2 `ncu --set full` captures (one GPU): the shapes VERDICT r01 asked about (1x1 256->1024 @14, 512->2048 @7, 3x3 64->64 @56),
# the two kernels bench.py's roofline objects quote, and the dgrad with the fused BatchNorm-backward sums.
# Raw page -> gpurun_out/r2ncu/*.csv (copy to profiles/).  Each program is run once WITHOUT ncu first.
OUT=gpurun_out/r2ncu
mkdir -p $OUT
cap() {  # name, kernel regex, jobs
  local name=$1 rx=$2 jobs=$3
  timeout 120 python tools/one_kernel.py --iters 2 --jobs "$jobs" > $OUT/$name.plain.log 2>&1 || { echo "$name: plain run failed"; tail -3 $OUT/$name.plain.log; return; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$rx" -o $OUT/$name -f python tools/one_kernel.py --iters 2 --jobs "$jobs" > $OUT/$name.ncu.log 2>&1
  echo "$name: ncu exit=$?"
  ncu -i $OUT/$name.ncu-rep --page raw --csv > $OUT/$name.csv 2>/dev/null
  rm -f $OUT/$name.ncu-rep
}
cap 1x1_256_1024_14 "conv_tc" "fprop:256,14,256,1024,1;dgrad:256,14,256,1024,1;wgrad_acc:256,14,256,1024,1"
cap 1x1_512_2048_7  "conv_tc" "fprop:256,7,512,2048,1;dgrad:256,7,512,2048,1;wgrad_acc:256,7,512,2048,1"
cap 3x3_64_56       "conv_tc" "fprop:256,56,64,64,3;dgrad:256,56,64,64,3;wgrad_acc:256,56,64,64,3"
cap roofline        "conv_tc|bn_bwd_apply" "fprop:256,14,256,256,3;bn_bwd_apply:1,1,8,8,1"
cap dgrad_bnb       "conv_tc" "dgrad:256,56,64,64,3;dgrad_bnb:256,56,64,64,3;dgrad:256,14,256,256,3;dgrad_bnb:256,14,256,256,3"
ls -la $OUT/*.csv
