#!/usr/bin/env bash
# GPU-box call: GPU tests on the default chain, bench lines default / LGK_TAIL_STAGE=0 / LGK_CHAIN=3, ncu launch list
set -u
TAG=${1:-r02z}
OUT=gpurun_out
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q --maxfail=30 > $OUT/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?"; tail -30 $OUT/pytest_gpu_$TAG.log | cut -c1-300
one() {  # name envs  (env vars from the caller)
  timeout 300 python bench.py --envs $2 --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>>$OUT/bench_$TAG.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('bench', '$1', $2, d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['launches_per_step'], d['scan_timeouts'])"
}
for rep in 1 2; do
  one default 65536
  LGK_TAIL_STAGE=0 one nostage 65536
  LGK_CHAIN=2w one wait 65536
  LGK_CHAIN=3 one chain3 65536
done
for v in ${VARIANTS:-}; do
  for rep in 1 2; do LGK_LIB_PATH=$PWD/build/liblgk_$v.so one $v 65536; done
done
for e in 4096 16384; do
  one default $e
  LGK_TAIL_STAGE=0 one nostage $e
  LGK_CHAIN=3 one chain3 $e
done
tail -5 $OUT/bench_$TAG.err
CMD="python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-e2e"
LGK_CHAIN=2w timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_launch_$TAG.log 2>&1
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
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1]))[:4]:
    print(f"{len(v):4d} x {sum(v) / len(v) / 1e3:9.2f} us  (min {min(v) / 1e3:8.2f})  {k}")
PY
