#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest -q -m gpu -p no:cacheprovider tests/test_gpu_a_ops.py tests/test_gpu_c_blocks.py tests/test_zz_no_alloc.py -x 2>&1 | tail -1
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-roofline"
run() { echo -n "$1: "; env $1 timeout 300 $B $2 2>/dev/null | python -c "import json,sys;d=json.loads(sys.stdin.read());print(d['ms_per_step'], d['value'])"; }
run "X=0"
run "X=0"
run "DV_BN_BWD_CTAS=8 DV_BN_APPLY_CTAS=8 DV_BN_RED_BLOCKS=1184"
run "X=0" "--model resnet34"
run "X=0" "--model mobilenet1"
run "X=0" "--model resnet152"
