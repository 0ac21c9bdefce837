#!/bin/bash
mkdir -p gpurun_out
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-roofline"
timeout 300 $B > gpurun_out/ab_none.json 2>/dev/null
DV_ATOMIC_WGRAD=1 timeout 300 $B > gpurun_out/ab_wgrad.json 2>/dev/null
DV_ATOMIC_BN=1 timeout 300 $B > gpurun_out/ab_bn.json 2>/dev/null
DV_DETERMINISTIC=1 timeout 300 $B > gpurun_out/ab_det.json 2>/dev/null
for f in none wgrad bn det; do echo -n "$f "; python -c "import json;d=json.load(open('gpurun_out/ab_$f.json'));print(d['ms_per_step'], d['value'], d['launches_per_step'])"; done
timeout 600 python -m pytest -q -m gpu -p no:cacheprovider tests -x > gpurun_out/all.log 2>&1; echo "tests exit=$?"; tail -1 gpurun_out/all.log
