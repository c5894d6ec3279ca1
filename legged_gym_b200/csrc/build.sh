#!/usr/bin/env bash
# Builds liblgk.so in-tree for sm_100a. Usage: build.sh [extra nvcc flags]
set -euo pipefail
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xptxas -v"
$NVCC $FLAGS -shared -o liblgk.so lgk_abi.cu lgk_scan.cu lgk_terrain.cu lgk_step.cu lgk_reset.cu lgk_control.cu lgk_lstm.cu lgk_mlp_tc.cu lgk_mlp_umma.cu "$@"
echo "built $(pwd)/liblgk.so"
