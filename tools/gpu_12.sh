#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/summary.txt
run() { name=$1; shift; timeout 600 python -m pytest -q -m gpu -p no:cacheprovider "$@" > gpurun_out/$name.log 2>&1; echo "$name exit=$?" >> gpurun_out/summary.txt; grep -E "passed|failed" gpurun_out/$name.log | tail -1 >> gpurun_out/summary.txt; }
run all tests
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-roofline"
timeout 300 $B > gpurun_out/bench_occ1.json 2> gpurun_out/bench_occ1.err
DV_TC_OCC=2 timeout 300 $B > gpurun_out/bench_occ2.json 2> /dev/null
timeout 300 python tools/layer_bench.py --what conv --out gpurun_out/lb_occ1 > /dev/null 2>&1
DV_TC_OCC=2 timeout 300 python tools/layer_bench.py --what conv --out gpurun_out/lb_occ2 > /dev/null 2>&1
timeout 300 $B --model mobilenet1 > gpurun_out/bench_mobilenet1.json 2>/dev/null
timeout 300 $B --model resnet34 > gpurun_out/bench_resnet34.json 2>/dev/null
timeout 300 $B --model resnet152 > gpurun_out/bench_resnet152.json 2>/dev/null
cat gpurun_out/summary.txt
for f in occ1 occ2 mobilenet1 resnet34 resnet152; do echo -n "$f "; python -c "import json;d=json.load(open('gpurun_out/bench_$f.json'));print(d['ms_per_step'], d['value'], d['launches_per_step'])"; done
grep -E "^E  |FAILED" gpurun_out/all.log | head -20
