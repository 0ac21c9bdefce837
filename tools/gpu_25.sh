#!/bin/bash
mkdir -p gpurun_out
for m in resnet34 mobilenet_v1; do
  timeout 300 python tools/step_trace.py --model $m --out gpurun_out/step_trace_$m.md > /dev/null 2>&1
  head -30 gpurun_out/step_trace_$m.md
done
timeout 300 python tools/step_trace.py --model resnet152 --batch 128 --out gpurun_out/step_trace_resnet152.md > /dev/null 2>&1
head -12 gpurun_out/step_trace_resnet152.md
