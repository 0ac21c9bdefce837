#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/summary.txt
run() { name=$1; shift; timeout 600 python -m pytest -q -m gpu -p no:cacheprovider "$@" > gpurun_out/$name.log 2>&1; echo "$name exit=$?" >> gpurun_out/summary.txt; grep -E "passed|failed" gpurun_out/$name.log | tail -1 >> gpurun_out/summary.txt; }
run b_tc tests/test_gpu_b_tc.py
timeout 300 python tools/layer_bench.py --what conv --out gpurun_out/lb_staged > /dev/null 2>&1
DV_TC_DIRECT=1 timeout 300 python tools/layer_bench.py --what conv --out gpurun_out/lb_direct > /dev/null 2>&1
LB_STATS=1 timeout 300 python tools/layer_bench.py --what conv --out gpurun_out/lb_staged_stats > /dev/null 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-roofline > gpurun_out/bench_a.json 2> gpurun_out/bench_a.err
DV_TC_DIRECT=1 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-roofline > gpurun_out/bench_direct_nofuse.json 2> gpurun_out/bench_b.err
LPS=$(python -c "import json;print(json.load(open('gpurun_out/bench_a.json'))['launches_per_step'])" 2>/dev/null || echo 545)
timeout 600 python bench.py --profile --no-graph --steps 2 --warmup 3 > gpurun_out/profile_plain.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s $((LPS*4+200)) -c $((LPS+160)) --csv --log-file gpurun_out/launches_r01_v4.csv \
    python bench.py --profile --no-graph --steps 2 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
echo "ncu exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; cat gpurun_out/bench_a.json gpurun_out/bench_direct_nofuse.json | cut -c1-200
