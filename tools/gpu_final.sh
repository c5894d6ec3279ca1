#!/usr/bin/env bash
# GPU-box call at the end of a round: GPU tests, one full bench line per BASELINE.json config + headline + reference
# arm, ncu launch list and ncu --set full of the two kernels of post_physics_step
set -u
TAG=${1:-r03}
OUT=gpurun_out
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.txt 2>&1
echo "pytest rc=$?"; tail -3 $OUT/${TAG}_pytest_gpu.txt | cut -c1-200
bash tools/gpu_bench_configs.sh $TAG
CMD="python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_launch.log 2>&1
echo "ncu launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'lgk_(tail|step)_kernel' -s 12 -c 2 -o $OUT/${TAG}_prof $CMD > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full rc=$?"; ls -la $OUT/${TAG}_prof.ncu-rep
