#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest -q -m gpu -p no:cacheprovider tests -x > gpurun_out/all.log 2>&1; echo "tests exit=$?"; tail -1 gpurun_out/all.log
grep -E "^E  |FAILED|timed out" gpurun_out/all.log | head
for m in resnet50 resnet34 mobilenet_v1; do timeout 200 python tools/infer_bench.py --model $m 2>&1 | tail -1; done | tee gpurun_out/infer_bench.jsonl
