#!/bin/bash
mkdir -p gpurun_out
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-roofline"
for rep in 1 2; do
DV_TC_DEBUG=1 timeout 300 $B > gpurun_out/ab_staged_fuse_$rep.json 2> gpurun_out/ab_staged_fuse_$rep.err
DV_TC_DIRECT=1 timeout 300 $B > gpurun_out/ab_direct_fuse_$rep.json 2>/dev/null
DV_FUSE_STATS=0 timeout 300 $B > gpurun_out/ab_staged_nofuse_$rep.json 2>/dev/null
DV_FUSE_STATS=0 DV_TC_DIRECT=1 timeout 300 $B > gpurun_out/ab_direct_nofuse_$rep.json 2>/dev/null
done
grep "dv conv_tc" gpurun_out/ab_staged_fuse_1.err
for f in gpurun_out/ab_*.json; do echo -n "$f "; python -c "import json,sys;d=json.load(open('$f'));print(d['ms_per_step'], d['value'])"; done
