#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/summary.txt
run() { name=$1; shift; timeout 600 python -m pytest -q -m gpu -p no:cacheprovider "$@" > gpurun_out/$name.log 2>&1; echo "$name exit=$?" >> gpurun_out/summary.txt; grep -E "passed|failed" gpurun_out/$name.log | tail -1 >> gpurun_out/summary.txt; }
run a_ops tests/test_gpu_a_ops.py
run b_tc tests/test_gpu_b_tc.py
run c_blocks tests/test_gpu_c_blocks.py
run d_models tests/test_gpu_d_models.py
run zz tests/test_zz_no_alloc.py
timeout 600 python tools/layer_bench.py --out gpurun_out/layer_bench_v3 > gpurun_out/layer_bench.log 2>&1; echo "layer_bench exit=$?" >> gpurun_out/summary.txt
timeout 600 python bench.py --steps 10 --warmup 3 --cpu-batch 16 > gpurun_out/bench_graph.json 2> gpurun_out/bench_graph.err; echo "bench_graph exit=$?" >> gpurun_out/summary.txt
timeout 600 python bench.py --model mobilenet1 --steps 10 --warmup 3 --no-cpu-baseline --no-roofline > gpurun_out/bench_mobilenet1.json 2> gpurun_out/bench_mobilenet1.err; echo "bench_mobilenet exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt; cat gpurun_out/bench_graph.json gpurun_out/bench_mobilenet1.json; tail -n 3 gpurun_out/bench_graph.err; grep -E "^E  |FAILED" gpurun_out/*.log | head -20
