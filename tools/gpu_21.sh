#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest -q -m gpu -p no:cacheprovider tests -x > gpurun_out/all.log 2>&1; echo "tests exit=$?"; tail -1 gpurun_out/all.log
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline"
timeout 300 $B > gpurun_out/bench_pair.json 2> gpurun_out/bench_pair.err
DV_TC_PAIR=0 timeout 300 $B > gpurun_out/bench_nopair.json 2> /dev/null
timeout 300 python tools/layer_bench.py --what conv --out gpurun_out/lb_lean > /dev/null 2>&1
DV_TC_PAIR=0 timeout 300 python tools/layer_bench.py --what conv --out gpurun_out/lb_lean_nopair > /dev/null 2>&1
for f in pair nopair; do echo -n "$f "; python -c "import json;d=json.load(open('gpurun_out/bench_$f.json'));print(d['ms_per_step'], d['value'], d['roofline'])"; done
grep -E "^E  |FAILED|timed out" gpurun_out/all.log | head
