#!/usr/bin/env bash
# GPU-box call: GPU tests on the default library, then bench lines + ncu launch lists for the library variants named
# in VARIANTS (build/liblgk_<name>.so; "default" = the in-tree library)
set -u
TAG=${1:-r02n}
OUT=gpurun_out
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q --maxfail=10 > $OUT/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?"; tail -6 $OUT/pytest_gpu_$TAG.log | cut -c1-300
for lib in ${VARIANTS:-default}; do
  p=$PWD/legged_gym_b200/csrc/liblgk.so; [ $lib != default ] && p=$PWD/build/liblgk_$lib.so
  for e in ${ENVS:-65536 4096}; do
  LGK_LIB_PATH=$p timeout 300 python bench.py --envs $e --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>>$OUT/bench_$TAG.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('bench', '$lib', $e, d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'])"
  done
  CMD="python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-e2e"
  LGK_LIB_PATH=$p timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $OUT/launches_${TAG}_$lib.csv $CMD > $OUT/ncu_launch_$TAG.log 2>&1
  python - <<PY
import csv, collections
rows = list(csv.reader(open("$OUT/launches_${TAG}_$lib.csv", errors="ignore")))
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
h = rows[hdr]; kn = h.index("Kernel Name"); mv = h.index("Metric Value")
agg = collections.defaultdict(list)
for r in rows[hdr + 1:]:
    if len(r) > mv:
        try: agg[r[kn][:70]].append(float(r[mv].replace(",", "")))
        except ValueError: pass
print("== $lib")
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1]))[:4]:
    print(f"{len(v):4d} x {sum(v) / len(v) / 1e3:9.2f} us  (min {min(v) / 1e3:8.2f})  {k}")
PY
done
