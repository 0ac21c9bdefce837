#!/bin/bash
# 2-GPU data-parallel bench (NCCL bucketed all-reduce): eager first (safe), then with CUDA-graph capture
mkdir -p gpurun_out
: > gpurun_out/summary.txt
nvidia-smi -L >> gpurun_out/summary.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 2 --steps 10 --warmup 3 --no-graph > gpurun_out/bench_dp2_nograph.json 2> gpurun_out/bench_dp2_nograph.err; echo "dp2_nograph exit=$?" >> gpurun_out/summary.txt
timeout 600 $TR bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_dp2_graph.json 2> gpurun_out/bench_dp2_graph.err; echo "dp2_graph exit=$?" >> gpurun_out/summary.txt
timeout 300 $TR bench.py --gpus 2 --impl reference --steps 1 --warmup 1 --cpu-batch 8 > gpurun_out/bench_dp2_ref.json 2> gpurun_out/bench_dp2_ref.err; echo "dp2_ref exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; cat gpurun_out/bench_dp2_nograph.json gpurun_out/bench_dp2_graph.json gpurun_out/bench_dp2_ref.json; tail -n 5 gpurun_out/bench_dp2_graph.err
