#!/usr/bin/env bash
# GPU-box call: smoke + GPU tests of the side-by-side scan / step kernels, then bench lines over LGK_SCAN_CTAS
set -u
TAG=${1:-r02d}
OUT=gpurun_out
mkdir -p $OUT
timeout 300 python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1
rc=$?; echo "smoke rc=$rc"; tail -4 $OUT/smoke_$TAG.log
if [ $rc -ne 0 ]; then echo "smoke failed: stopping"; exit 1; fi
timeout 1200 python -m pytest tests -m gpu -q --maxfail=25 > $OUT/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?"; tail -30 $OUT/pytest_gpu_$TAG.log | cut -c1-300
for s in 4 2 3 5 6; do
  echo "LGK_SCAN_CTAS=$s"
  LGK_SCAN_CTAS=$s timeout 300 python bench.py --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>>$OUT/bench_$TAG.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'])"
done
for s in 4 3; do
  echo "4096 envs LGK_SCAN_CTAS=$s"
  LGK_SCAN_CTAS=$s timeout 300 python bench.py --envs 4096 --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>>$OUT/bench_$TAG.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'])"
done
tail -5 $OUT/bench_$TAG.err
