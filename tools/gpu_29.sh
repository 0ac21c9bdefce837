#!/bin/bash
mkdir -p gpurun_out
DV_SIDE_WGRAD=1 timeout 600 python -m pytest -q -m gpu -p no:cacheprovider tests/test_gpu_e_trainer.py -x > gpurun_out/side.log 2>&1; echo "trainer tests (side) exit=$?"; tail -1 gpurun_out/side.log
grep -E "^E  |FAILED|timed out" gpurun_out/side.log | head
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-roofline"
for v in 0 1 0 1; do
  DV_SIDE_WGRAD=$v timeout 300 $B > gpurun_out/bench_side$v.json 2> gpurun_out/bench_side$v.err
  echo -n "side=$v "; python -c "import json;d=json.load(open('gpurun_out/bench_side$v.json'));print(d['ms_per_step'], d['value'])"
done
for m in resnet34 resnet152 mobilenet1; do for v in 0 1; do
  DV_SIDE_WGRAD=$v timeout 300 $B --model $m > gpurun_out/bench_s.json 2> /dev/null
  echo -n "$m side=$v "; python -c "import json;d=json.load(open('gpurun_out/bench_s.json'));print(d['ms_per_step'], d['value'])"
done; done
