#!/bin/bash
# Round-2 multi-GPU run on ONE box with N GPUs (default 8): BASELINE configs[2] (ResNet-50), [3] (MobileNetV1), [4] (ResNet-152 b128),
# the two-rank NCCL tests, and a CUPTI timeline of rank 0 with the all-reduce kernels in it.
N=${1:-8}
OUT=gpurun_out/r2scale
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_g_multi.py -q > $OUT/multi_test.log 2>&1; echo "multi tests exit=$?"; tail -2 $OUT/multi_test.log
for m in resnet50 mobilenet1 resnet152; do
  timeout 400 $TR --master-port 29511 bench.py --gpus $N --model $m --steps 20 --warmup 5 > $OUT/bench_${m}_dp$N.json 2> $OUT/bench_${m}_dp$N.err; echo "$m dp$N exit=$?"
  cut -c1-330 $OUT/bench_${m}_dp$N.json
done
timeout 300 $TR --master-port 29512 tools/step_trace.py --model resnet50 --out $OUT/step_trace_n$N.md > $OUT/trace.log 2>&1; echo "trace exit=$?"
grep -i "nccl\|span per step" $OUT/step_trace_n$N.md | head
timeout 300 $TR --master-port 29513 tools/step_trace.py --model mobilenet1 --out $OUT/step_trace_mobilenet1_n$N.md > $OUT/trace2.log 2>&1
grep -i "nccl\|span per step" $OUT/step_trace_mobilenet1_n$N.md | head
for m in resnet50; do timeout 300 python bench.py --model $m --steps 20 --warmup 5 --no-cpu-baseline --no-roofline > $OUT/bench_${m}_dp1.json 2>/dev/null; cut -c1-200 $OUT/bench_${m}_dp1.json; done
