#!/usr/bin/env bash
# GPU-box call: A/B of the staged-terrain-tile scan kernel (build/liblgk_staged.so, -DLGK_SCAN_STAGED) against the default
set -u
OUT=gpurun_out; mkdir -p $OUT
LGK_LIB_PATH=$PWD/build/liblgk_staged.so timeout 900 python -m pytest tests -m gpu -q -x > $OUT/pytest_gpu_staged.log 2>&1
echo "pytest(staged) rc=$?"; tail -4 $OUT/pytest_gpu_staged.log | cut -c1-300
for lib in default staged; do
  p=$PWD/legged_gym_b200/csrc/liblgk.so; [ $lib != default ] && p=$PWD/build/liblgk_$lib.so
  for r in 1 2; do
  LGK_LIB_PATH=$p timeout 300 python bench.py --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>>$OUT/bench_staged.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('bench', '$lib', d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d.get('scan_timeouts'))"
  done
  CMD="python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-e2e"
  LGK_LIB_PATH=$p timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum --clock-control none -k regex:lgk_scan -c 12 --csv --log-file $OUT/scan_ab_$lib.csv $CMD > $OUT/ncu_staged.log 2>&1
  python - <<PY
import csv, collections
rows = list(csv.reader(open("$OUT/scan_ab_$lib.csv", errors="ignore")))
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
h = rows[hdr]; mn = h.index("Metric Name"); mv = h.index("Metric Value")
agg = collections.defaultdict(list)
for r in rows[hdr + 1:]:
    if len(r) > mv:
        try: agg[r[mn]].append(float(r[mv].replace(",", "")))
        except ValueError: pass
print("== $lib", {k: round(sum(v) / len(v), 1) for k, v in agg.items()})
PY
done
