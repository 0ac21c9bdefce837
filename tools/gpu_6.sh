#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/summary.txt
run() { name=$1; shift; timeout 900 python -m pytest -q -m gpu -p no:cacheprovider "$@" > gpurun_out/$name.log 2>&1; echo "$name exit=$?" >> gpurun_out/summary.txt; grep -E "passed|failed" gpurun_out/$name.log | tail -1 >> gpurun_out/summary.txt; }
run a_ops tests/test_gpu_a_ops.py
run b_tc tests/test_gpu_b_tc.py
run c_blocks tests/test_gpu_c_blocks.py
run d_models tests/test_gpu_d_models.py
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit=$?" >> gpurun_out/summary.txt
timeout 900 python bench.py --steps 10 --warmup 3 --cpu-batch 16 > gpurun_out/bench_graph.json 2> gpurun_out/bench_graph.err; echo "bench_graph exit=$?" >> gpurun_out/summary.txt
LPS=$(python -c "import json;print(json.load(open('gpurun_out/bench_graph.json'))['launches_per_step'])" 2>/dev/null || echo 600)
timeout 600 python bench.py --profile --no-graph --steps 2 --warmup 3 > gpurun_out/profile_plain.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s $((LPS*4+30)) -c $((LPS)) --csv --log-file gpurun_out/launches_r01_v3.csv \
    python bench.py --profile --no-graph --steps 2 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
echo "ncu exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; cat gpurun_out/bench_graph.json; tail -n 3 gpurun_out/bench_graph.err; grep -E "^E|FAILED" gpurun_out/*.log | head -20
