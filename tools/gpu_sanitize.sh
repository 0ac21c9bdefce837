#!/bin/bash
# compute-sanitizer over the kernel unit tests (small shapes) and one eager Trainer step with the side-stream wgrad on.
# Logs -> gpurun_out/sanitizer/ (copy the summaries to profiles/).  Each run is bounded by `timeout`.
OUT=gpurun_out/sanitizer
mkdir -p $OUT
CS=/usr/local/cuda/bin/compute-sanitizer
run() {  # name, tool, command...
  local name=$1 tool=$2; shift 2
  timeout ${SAN_TIMEOUT:-1200} $CS --tool $tool --error-exitcode 77 "$@" > $OUT/$name.log 2>&1
  echo "$name: exit=$? $(grep 'ERROR SUMMARY' $OUT/$name.log | tail -1)"
  tail -2 $OUT/$name.log
}
SMALL='oracle and not 151x and not 149x and not 38x'
for t in ${SAN_TOOLS:-memcheck racecheck synccheck}; do
  run ${t}_tc   $t python -m pytest tests/test_gpu_b_tc.py -x -q -k "$SMALL"
  run ${t}_ops  $t python -m pytest tests/test_gpu_a_ops.py -x -q
  run ${t}_trainer $t python tools/grad_noise.py --model resnet34 --batch 2 --size 64 --out $OUT/grad_noise_$t.md
done
grep -H "ERROR SUMMARY" $OUT/*.log
