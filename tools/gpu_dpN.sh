#!/bin/bash
# usage: gpu_dpN.sh N  — data-parallel bench at N GPUs (our arm + reference arm), as the driver launches it
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_dp$N.json 2> gpurun_out/bench_dp$N.err; echo "dp$N exit=$?"
timeout 200 $TR bench.py --gpus $N --impl reference --steps 1 --warmup 1 --cpu-batch 8 > gpurun_out/bench_dp${N}_ref.json 2> gpurun_out/bench_dp${N}_ref.err; echo "dp${N}_ref exit=$?"
cat gpurun_out/bench_dp$N.json | cut -c1-400; tail -n 3 gpurun_out/bench_dp$N.err | cut -c1-300
