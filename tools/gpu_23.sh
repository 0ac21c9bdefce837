#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest -q -m gpu -p no:cacheprovider tests -x > gpurun_out/all.log 2>&1; echo "tests exit=$?"; tail -1 gpurun_out/all.log
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-roofline"
timeout 300 $B > gpurun_out/bench_a.json 2> gpurun_out/bench_a.err
for f in a; do echo -n "$f "; python -c "import json;d=json.load(open('gpurun_out/bench_$f.json'));print(d['ms_per_step'], d['value'], d['e2e'])"; done
grep -E "^E  |FAILED|timed out" gpurun_out/all.log | head
