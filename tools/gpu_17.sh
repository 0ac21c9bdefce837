#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest -q -m gpu -p no:cacheprovider tests -x > gpurun_out/all.log 2>&1; echo "tests exit=$?"; tail -2 gpurun_out/all.log
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-roofline"
timeout 300 $B > gpurun_out/bench_v7.json 2> gpurun_out/bench_v7.err
DV_DETERMINISTIC=1 timeout 300 $B > gpurun_out/bench_v7_det.json 2> /dev/null
timeout 300 python tools/layer_bench.py --what bn --out gpurun_out/lb_v7 > /dev/null 2>&1
timeout 300 $B --model mobilenet1 > gpurun_out/bench_mobilenet1.json 2>/dev/null
timeout 300 $B --model resnet34 > gpurun_out/bench_resnet34.json 2>/dev/null
timeout 300 $B --model resnet152 > gpurun_out/bench_resnet152.json 2>/dev/null
for f in v7 v7_det mobilenet1 resnet34 resnet152; do echo -n "$f "; python -c "import json;d=json.load(open('gpurun_out/bench_$f.json'));print(d['ms_per_step'], d['value'], d['launches_per_step'])"; done
grep -E "^E  |FAILED" gpurun_out/all.log | head
