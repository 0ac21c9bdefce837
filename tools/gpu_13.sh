#!/bin/bash
mkdir -p gpurun_out
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-roofline"
timeout 300 python tools/layer_bench.py --what conv --out gpurun_out/lb_v5 > /dev/null 2>&1
timeout 300 $B > gpurun_out/bench_v5.json 2> gpurun_out/bench_v5.err
DV_TC_OCC=1 timeout 300 $B > gpurun_out/bench_v5_occ1.json 2> /dev/null
timeout 300 python -m pytest -q -m gpu -p no:cacheprovider tests/test_gpu_b_tc.py > gpurun_out/b_tc.log 2>&1; tail -1 gpurun_out/b_tc.log
for f in v5 v5_occ1; do echo -n "$f "; python -c "import json;d=json.load(open('gpurun_out/bench_$f.json'));print(d['ms_per_step'], d['value'], d['launches_per_step'])"; done
