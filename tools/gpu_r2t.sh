#!/usr/bin/env bash
# GPU-box call: GPU tests on the fused two-launch chain (default) and on the three-launch chain (LGK_CHAIN=3), bench
# lines for both and for library variants (build/liblgk_<name>.so), ncu launch list of the default
set -u
TAG=${1:-r02t}
OUT=gpurun_out
mkdir -p $OUT
timeout 300 python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1
echo "smoke rc=$?"; tail -3 $OUT/smoke_$TAG.log | cut -c1-300
for ch in 2w 2f; do
LGK_CHAIN=$ch timeout 1500 python -m pytest tests -m gpu -q --maxfail=30 > $OUT/pytest_gpu_${TAG}_$ch.log 2>&1
echo "pytest $ch rc=$?"; tail -30 $OUT/pytest_gpu_${TAG}_$ch.log | cut -c1-300
done
one() {  # name chain lib envs
  local p=$PWD/legged_gym_b200/csrc/liblgk.so; [ $3 != default ] && p=$PWD/build/liblgk_$3.so
  LGK_CHAIN=$2 LGK_LIB_PATH=$p timeout 300 python bench.py --envs $4 --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>>$OUT/bench_$TAG.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('bench', '$1', $4, d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['launches_per_step'], d['scan_timeouts'])"
}
for rep in 1 2; do
  one fused_wait 2w default 65536
  one fused_flags 2f default 65536
  one chain3 3 default 65536
done
for v in ${VARIANTS:-}; do one ${v}_wait 2w $v 65536; one ${v}_flags 2f $v 65536; done
for e in 4096 16384; do
one fused_wait 2w default $e
one fused_flags 2f default $e
one chain3 3 default $e
done
tail -5 $OUT/bench_$TAG.err
CMD="python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-e2e"
LGK_CHAIN=${NCU_CHAIN:-2w} timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_launch_$TAG.log 2>&1
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
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1]))[:5]:
    print(f"{len(v):4d} x {sum(v) / len(v) / 1e3:9.2f} us  (min {min(v) / 1e3:8.2f})  {k}")
PY
