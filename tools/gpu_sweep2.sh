#!/bin/bash
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-roofline"
run() { echo -n "$1 $2: "; env $1 timeout 300 $B $2 2>/dev/null | python -c "import json,sys;d=json.loads(sys.stdin.read());print(d['ms_per_step'], d['value'])"; }
for c in 16 2 3 4 6; do run "DV_DW_FWD_CAP=$c" "--model mobilenet1"; done
for c in 2 3 4; do run "DV_DW_S2_CAP=$c" "--model mobilenet1"; done
for c in 8 2 3 4; do run "DV_POOL_CAP=$c" "--model resnet34"; done
