#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/summary.txt
LPS=597
timeout 600 python bench.py --profile --no-graph --steps 1 --warmup 3 > gpurun_out/profile_plain.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s $((LPS*4)) -c $((LPS+10)) --csv --log-file gpurun_out/launches_r01_v2.csv \
    python bench.py --profile --no-graph --steps 1 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
echo "ncu exit=$?" >> gpurun_out/summary.txt
for m in resnet34 resnet152 mobilenet1; do
timeout 600 python bench.py --model $m --steps 10 --warmup 3 --no-cpu-baseline --no-roofline > gpurun_out/bench_$m.json 2> gpurun_out/bench_$m.err; echo "bench_$m exit=$?" >> gpurun_out/summary.txt
done
cat gpurun_out/summary.txt; cat gpurun_out/bench_resnet34.json gpurun_out/bench_resnet152.json gpurun_out/bench_mobilenet1.json
