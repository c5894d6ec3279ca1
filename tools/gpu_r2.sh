#!/usr/bin/env bash
# One GPU-box call of round 2: smoke, GPU tests, bench lines, ncu launch list (+ optional full capture).
# usage (here): gpurun --timeout 1500 -- 'bash tools/gpu_r2.sh <tag> [full]'
set -u
TAG=${1:-r02a}
FULL=${2:-}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm,power.draw --format=csv | tee $OUT/smi_$TAG.txt
# a hang must not eat the box: the first contact with the new kernels runs under a short timeout
timeout 300 python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1
rc=$?; echo "smoke rc=$rc"; tail -4 $OUT/smoke_$TAG.log
if [ $rc -ne 0 ]; then echo "smoke failed: stopping"; exit 1; fi
timeout 1200 python -m pytest tests -m gpu -q --maxfail=25 > $OUT/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?"; tail -40 $OUT/pytest_gpu_$TAG.log | cut -c1-300
timeout 600 python bench.py --no-cpu-baseline > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err
echo "bench rc=$?"; cat $OUT/bench_$TAG.json; tail -3 $OUT/bench_$TAG.err
timeout 300 python bench.py --envs 4096 --no-cpu-baseline > $OUT/bench4096_$TAG.json 2>> $OUT/bench_$TAG.err
cat $OUT/bench4096_$TAG.json
CMD="python bench.py --steps 8 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > $OUT/plain_$TAG.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_launch_$TAG.log 2>&1
echo "ncu launches rc=$?"
python - <<EOF
import csv, collections
rows = list(csv.reader(open("$OUT/launches_$TAG.csv", errors="ignore")))
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
h = rows[hdr]; kn = h.index("Kernel Name"); mv = h.index("Metric Value")
agg = collections.defaultdict(list)
for r in rows[hdr + 1:]:
    if len(r) > mv:
        try: agg[r[kn][:70]].append(float(r[mv].replace(",", "")))
        except ValueError: pass
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print(f"{len(v):4d} x {sum(v) / len(v) / 1e3:9.2f} us  (min {min(v) / 1e3:8.2f})  {k}")
EOF
if [ -n "$FULL" ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"lgk_scan_kernel|lgk_step_kernel" -s 12 -c 2 -o $OUT/prof_$TAG $CMD > $OUT/ncu_full_$TAG.log 2>&1
  echo "ncu full rc=$?"
fi
ls -la $OUT | tail -15
