#!/usr/bin/env bash
# ncu --set full capture of the step kernel only. usage: gpurun -- 'bash tools/gpu_prof.sh <tag>'
set -u
TAG=${1:-x}
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --steps 8 --warmup 3 --no-cpu-baseline"
$CMD > $OUT/plain_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lgk_step_kernel -s 12 -c 1 -o $OUT/prof_step_$TAG $CMD > $OUT/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"
tail -1 $OUT/plain_$TAG.log | cut -c1-200
