#!/bin/bash
# first GPU bring-up: each group in its own process so one faulting kernel does not hide the others
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version --format=csv > gpurun_out/gpu.txt 2>&1
run() { name=$1; shift; timeout 600 python -m pytest -q -m gpu -p no:cacheprovider "$@" > gpurun_out/$name.log 2>&1; echo "$name exit=$?" >> gpurun_out/summary.txt; tail -3 gpurun_out/$name.log | head -2 >> gpurun_out/summary.txt; }
: > gpurun_out/summary.txt
run a_ops tests/test_gpu_a_ops.py
run tc_fprop_1x1 tests/test_gpu_b_tc.py -k "fprop and (2x64x8x8x64x1x1x0 or 3x128x7x7x256x1x1x0 or 2x256x8x8x128x1x1x0 or 2x512x4x4x1000x1x1x0)"
run tc_fprop_im2col tests/test_gpu_b_tc.py -k "fprop and not (2x64x8x8x64x1x1x0 or 3x128x7x7x256x1x1x0 or 2x256x8x8x128x1x1x0 or 2x512x4x4x1000x1x1x0)"
run tc_dgrad tests/test_gpu_b_tc.py -k "dgrad"
run tc_wgrad tests/test_gpu_b_tc.py -k "wgrad"
run tc_full tests/test_gpu_b_tc.py -k "full_size"
run c_blocks tests/test_gpu_c_blocks.py
run d_models tests/test_gpu_d_models.py
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
