#!/usr/bin/env bash
# GPU-box call: the allocation-free entry point's test alone in a fresh process, then the round-2 script
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/noalloc_alone.log 2>&1
echo "noalloc alone rc=$?"; tail -15 gpurun_out/noalloc_alone.log | cut -c1-250
bash tools/gpu_r2.sh "$@"
