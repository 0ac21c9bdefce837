#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest -q -m gpu -p no:cacheprovider tests -x > gpurun_out/all.log 2>&1; echo "tests exit=$?"; tail -1 gpurun_out/all.log
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-roofline"
timeout 300 $B > gpurun_out/bench_v8.json 2> gpurun_out/bench_v8.err
for f in v8; do echo -n "$f "; python -c "import json;d=json.load(open('gpurun_out/bench_$f.json'));print(d['ms_per_step'], d['value'], d['launches_per_step'])"; done
LPS=$(python -c "import json;print(json.load(open('gpurun_out/bench_v8.json'))['launches_per_step'])" 2>/dev/null || echo 440)
timeout 600 python bench.py --profile --no-graph --steps 2 --warmup 3 > gpurun_out/profile_plain.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s $((LPS*4+100)) -c $((LPS+30)) --csv --log-file gpurun_out/launches_r01_v8.csv \
    python bench.py --profile --no-graph --steps 2 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
echo "ncu exit=$?"
grep -E "^E  |FAILED" gpurun_out/all.log | head
