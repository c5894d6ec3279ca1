#!/usr/bin/env bash
# Step-kernel residency sweep on the GPU box: rebuilds liblgk.so with -DLGK_MIN_CTAS=<k> and times the step.
# usage: gpurun -- 'bash tools/sweep_min_ctas.sh 5 6 7'
set -u
mkdir -p gpurun_out
for k in "$@"; do
  (cd legged_gym_b200/csrc && bash build.sh -DLGK_MIN_CTAS=$k > /tmp/build_$k.log 2>&1; grep -A2 "step_kernelILi12ELb0ELb1ELi6" /tmp/build_$k.log | grep -E "spill|Used" | head -2)
  python bench.py --steps 200 --warmup 20 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('MIN_CTAS=$k', 'ms_per_step', d['ms_per_step'], 'kernel_ms', d['roofline']['kernel_ms'], 'frac', d['roofline']['frac'])"
done
(cd legged_gym_b200/csrc && bash build.sh > /dev/null 2>&1)
