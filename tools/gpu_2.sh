#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/summary.txt
run() { name=$1; shift; timeout 900 python -m pytest -q -m gpu -p no:cacheprovider "$@" > gpurun_out/$name.log 2>&1; echo "$name exit=$?" >> gpurun_out/summary.txt; grep -E "passed|failed" gpurun_out/$name.log | tail -1 >> gpurun_out/summary.txt; }
run c_blocks tests/test_gpu_c_blocks.py
run d_models tests/test_gpu_d_models.py -s
timeout 900 python bench.py --steps 5 --warmup 3 --no-graph --no-cpu-baseline > gpurun_out/bench_nograph.json 2> gpurun_out/bench_nograph.err; echo "bench_nograph exit=$?" >> gpurun_out/summary.txt
timeout 900 python bench.py --steps 10 --warmup 3 --cpu-batch 16 > gpurun_out/bench_graph.json 2> gpurun_out/bench_graph.err; echo "bench_graph exit=$?" >> gpurun_out/summary.txt
LPS=$(python -c "import json;print(json.load(open('gpurun_out/bench_nograph.json'))['launches_per_step'])" 2>/dev/null || echo 700)
echo "launches_per_step=$LPS" >> gpurun_out/summary.txt
# launch list of ONE eager step (the 5th: 1 counting step + 3 warm-up before it)
timeout 600 python bench.py --profile --no-graph --steps 1 --warmup 3 > gpurun_out/profile_plain.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s $((LPS*4)) -c $((LPS+50)) --csv --log-file gpurun_out/launches_r01.csv \
    python bench.py --profile --no-graph --steps 1 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
echo "ncu exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; cat gpurun_out/bench_nograph.json gpurun_out/bench_graph.json; tail -5 gpurun_out/bench_graph.err
