#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest -q -m gpu -p no:cacheprovider tests -x > gpurun_out/all.log 2>&1; echo "tests exit=$?"; tail -1 gpurun_out/all.log
grep -E "^E  |FAILED|timed out" gpurun_out/all.log | head
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-roofline"
for m in resnet34 resnet50 mobilenet_v1; do
  timeout 300 $B --model $m > gpurun_out/bench_$m.json 2> gpurun_out/bench_$m.err
  echo -n "$m "; python -c "import json;d=json.load(open('gpurun_out/bench_$m.json'));print(d['ms_per_step'], d['value'], d['e2e']['value'])"
done
timeout 300 $B --model resnet152 > gpurun_out/bench_resnet152.json 2> gpurun_out/bench_resnet152.err
echo -n "resnet152 "; python -c "import json;d=json.load(open('gpurun_out/bench_resnet152.json'));print(d['ms_per_step'], d['value'], d['config'])"
timeout 300 python tools/step_trace.py --model resnet34 --out gpurun_out/step_trace_resnet34.md > /dev/null 2>&1; head -16 gpurun_out/step_trace_resnet34.md
timeout 300 python tools/step_trace.py --model mobilenet_v1 --out gpurun_out/step_trace_mobilenet_v1.md 2>&1 | tail -5; head -24 gpurun_out/step_trace_mobilenet_v1.md
