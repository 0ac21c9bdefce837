#!/bin/bash
# Round-2 closing captures on one GPU: `ncu --set full` of the two kernels bench.py's roofline objects quote and of the 3x3 64->64
# halo-tile kernel (each program run once WITHOUT ncu first), CUPTI step timelines, one bench line per model.
OUT=gpurun_out/r2final
mkdir -p $OUT
cap() {  # name, kernel regex, jobs
  local name=$1 rx=$2 jobs=$3
  timeout 120 python tools/one_kernel.py --iters 2 --jobs "$jobs" > $OUT/$name.plain.log 2>&1 || { echo "$name: plain run failed"; tail -3 $OUT/$name.plain.log; return; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$rx" -o $OUT/$name -f python tools/one_kernel.py --iters 2 --jobs "$jobs" > $OUT/$name.ncu.log 2>&1
  echo "$name: ncu exit=$?"
  ncu -i $OUT/$name.ncu-rep --page raw --csv > $OUT/$name.csv 2>/dev/null
  rm -f $OUT/$name.ncu-rep
}
cap roofline   "conv_tc|bn_bwd_apply" "fprop:256,14,256,256,3;bn_bwd_apply:1,1,8,8,1"
cap 3x3_64_56_halo "conv" "fprop:256,56,64,64,3;fprop_stats:256,56,64,64,3;dgrad:256,56,64,64,3;wgrad_acc:256,56,64,64,3"
for m in resnet50 mobilenet1; do python tools/step_trace.py --model $m --out $OUT/trace_$m.md > $OUT/trace_$m.log 2>&1; done
python bench.py --steps 30 --warmup 5 > $OUT/bench_resnet50.json 2> $OUT/bench_resnet50.err
python bench.py --model resnet34 --steps 30 --warmup 5 > $OUT/bench_resnet34.json 2> $OUT/bench_resnet34.err
python bench.py --model mobilenet1 --steps 30 --warmup 5 > $OUT/bench_mobilenet1.json 2> $OUT/bench_mobilenet1.err
python bench.py --model resnet152 --batch 128 --steps 30 --warmup 5 > $OUT/bench_resnet152.json 2> $OUT/bench_resnet152.err
python bench.py --deterministic --steps 20 --warmup 5 --no-uint8 > $OUT/bench_deterministic.json 2> $OUT/bench_deterministic.err
for f in $OUT/bench_*.json; do python -c "import json,sys; d=json.loads(open('$f').readline()); print('$f', d['ms_per_step'], d['value'], d['e2e']['value'], d.get('roofline',{}).get('frac'))"; done
ls -la $OUT/*.csv
