#!/usr/bin/env bash
# ncu --set full capture of the two kernels of the fused chain (one launch each). usage: gpurun -- 'bash tools/gpu_prof2.sh <tag> [chain]'
set -u
TAG=${1:-x}; export LGK_CHAIN=${2:-2w}
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-e2e"
$CMD > $OUT/plain_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'lgk_(tail|step)_kernel' -s 12 -c 2 -o $OUT/prof_$TAG $CMD > $OUT/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"; ls -la $OUT/prof_$TAG.ncu-rep
tail -1 $OUT/plain_$TAG.log | cut -c1-200
