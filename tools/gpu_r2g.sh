#!/usr/bin/env bash
# GPU-box call: smoke, GPU tests, phase timings, bench lines, residency sweep of the step kernel, ncu launch list
set -u
TAG=${1:-r02g}
OUT=gpurun_out
mkdir -p $OUT
timeout 300 python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1
rc=$?; echo "smoke rc=$rc"; tail -4 $OUT/smoke_$TAG.log
if [ $rc -ne 0 ]; then echo "smoke failed: stopping"; exit 1; fi
timeout 1500 python -m pytest tests -m gpu -q --maxfail=25 > $OUT/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?"; tail -30 $OUT/pytest_gpu_$TAG.log | cut -c1-300
timeout 300 python tools/phase_times2.py 65536 2>&1 | tail -12
for e in 65536 16384 4096; do
timeout 300 python bench.py --envs $e --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>>$OUT/bench_$TAG.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('bench', $e, d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'])"
done
for m in 7 8; do
LGK_LIB_PATH=$PWD/build/liblgk_c$m.so timeout 300 python bench.py --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>>$OUT/bench_$TAG.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('bench min_ctas', $m, d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'])"
done
CMD="python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_launch_$TAG.log 2>&1
python - <<PY
import csv, collections
rows = list(csv.reader(open("$OUT/launches_$TAG.csv", errors="ignore")))
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
h = rows[hdr]; kn = h.index("Kernel Name"); mv = h.index("Metric Value")
agg = collections.defaultdict(list)
for r in rows[hdr + 1:]:
    if len(r) > mv:
        try: agg[r[kn][:70]].append(float(r[mv].replace(",", "")))
        except ValueError: pass
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1]))[:8]:
    print(f"{len(v):4d} x {sum(v) / len(v) / 1e3:9.2f} us  (min {min(v) / 1e3:8.2f})  {k}")
PY
if [ -n "${2:-}" ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"lgk_scan_kernel|lgk_step_kernel|lgk_finish_kernel" -s 18 -c 3 -o $OUT/prof_$TAG $CMD > $OUT/ncu_full_$TAG.log 2>&1
  echo "ncu full rc=$?"
fi
