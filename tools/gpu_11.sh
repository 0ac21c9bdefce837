#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/summary.txt
run() { name=$1; shift; timeout 600 python -m pytest -q -m gpu -p no:cacheprovider "$@" > gpurun_out/$name.log 2>&1; echo "$name exit=$?" >> gpurun_out/summary.txt; grep -E "passed|failed" gpurun_out/$name.log | tail -1 >> gpurun_out/summary.txt; }
run a_ops tests/test_gpu_a_ops.py
run b_tc tests/test_gpu_b_tc.py
run c_blocks tests/test_gpu_c_blocks.py
run d_models tests/test_gpu_d_models.py
run e_trainer tests/test_gpu_e_trainer.py
run zz tests/test_zz_no_alloc.py
DV_TC_DEBUG=1 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-roofline > gpurun_out/bench_a.json 2> gpurun_out/bench_a.err
grep "dv conv_tc" gpurun_out/bench_a.err >> gpurun_out/summary.txt
timeout 600 python bench.py --model mobilenet1 --steps 20 --warmup 3 --no-cpu-baseline --no-roofline > gpurun_out/bench_mobilenet1.json 2> gpurun_out/bench_mobilenet1.err
timeout 300 python tools/layer_bench.py --what conv,dw --out gpurun_out/layer_bench_v4 > /dev/null 2>&1
cat gpurun_out/summary.txt; cat gpurun_out/bench_a.json gpurun_out/bench_mobilenet1.json | cut -c1-220; grep -E "^E  |FAILED" gpurun_out/*.log | head -20
