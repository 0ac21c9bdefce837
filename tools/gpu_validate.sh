#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest -q -m gpu -p no:cacheprovider tests -x > gpurun_out/all.log 2>&1; echo "tests exit=$?"; tail -1 gpurun_out/all.log
grep -E "^E  |FAILED|timed out" gpurun_out/all.log | head
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench exit=$?"
cat gpurun_out/bench_default.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref exit=$?"
cat gpurun_out/bench_reference.json
LPS=$(python -c "import json;print(json.load(open('gpurun_out/bench_default.json'))['launches_per_step'])" 2>/dev/null || echo 460)
timeout 600 python bench.py --profile --no-graph --steps 2 --warmup 3 > gpurun_out/profile_plain.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s $((LPS*4+100)) -c $((LPS+30)) --csv --log-file gpurun_out/launches_r01_v10.csv \
    python bench.py --profile --no-graph --steps 2 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
echo "ncu exit=$?"
