#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest -q -m gpu -p no:cacheprovider tests -x > gpurun_out/all.log 2>&1; echo "tests exit=$?"; tail -2 gpurun_out/all.log
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-roofline"
timeout 300 python tools/layer_bench.py --what conv --out gpurun_out/lb_v6 > /dev/null 2>&1
timeout 300 $B > gpurun_out/bench_v6.json 2> gpurun_out/bench_v6.err
LB_STATS=1 timeout 300 python tools/layer_bench.py --what conv --out gpurun_out/lb_v6_stats > /dev/null 2>&1
for f in v6; do echo -n "$f "; python -c "import json;d=json.load(open('gpurun_out/bench_$f.json'));print(d['ms_per_step'], d['value'], d['launches_per_step'])"; done
grep -E "^E  |FAILED" gpurun_out/all.log | head
