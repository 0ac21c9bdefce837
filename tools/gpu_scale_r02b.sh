#!/bin/bash
# Closing multi-GPU lines of round 2 (one box, N GPUs): BASELINE configs[2..4] under torchrun + the 1-GPU ResNet-50 line of the same box.
N=${1:-8}
OUT=gpurun_out/r2scale_b
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for m in resnet50 mobilenet1 resnet152; do
  timeout 300 $TR --master-port 29511 bench.py --gpus $N --model $m --steps 20 --warmup 5 --no-cpu-baseline --no-roofline > $OUT/bench_${m}_dp$N.json 2> $OUT/bench_${m}_dp$N.err; echo "$m dp$N exit=$?"
  cut -c1-400 $OUT/bench_${m}_dp$N.json
done
timeout 200 python bench.py --model resnet50 --steps 20 --warmup 5 --no-cpu-baseline --no-roofline > $OUT/bench_resnet50_dp1.json 2>/dev/null; cut -c1-200 $OUT/bench_resnet50_dp1.json
