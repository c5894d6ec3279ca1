#!/usr/bin/env bash
# GPU-box call: the GPU tests with each chain variant forced (three launches / grid-wide wait / hand-over flags), smoke
set -u
TAG=${1:-r03}
OUT=gpurun_out; mkdir -p $OUT
timeout 300 python __graft_entry__.py smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/${TAG}_smoke.log | cut -c1-200
for ch in 3 2w 2f; do
  LGK_CHAIN=$ch timeout 1500 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu_chain_$ch.txt 2>&1
  echo "pytest LGK_CHAIN=$ch rc=$?"; tail -1 $OUT/${TAG}_pytest_gpu_chain_$ch.txt | cut -c1-200
done
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $OUT/${TAG}_noalloc_alone.txt 2>&1; echo "no-alloc alone rc=$?"; tail -1 $OUT/${TAG}_noalloc_alone.txt
