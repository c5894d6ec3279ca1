#!/usr/bin/env bash
# GPU-box call: GPU tests (default and LGK_TAIL_EPG=8), bench lines over the tail kernel's unit size
set -u
TAG=${1:-r03c}
OUT=gpurun_out
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q --maxfail=30 > $OUT/pytest_gpu_$TAG.log 2>&1
echo "pytest rc=$?"; tail -30 $OUT/pytest_gpu_$TAG.log | cut -c1-300
LGK_TAIL_EPG=${EPG2:-8} timeout 1500 python -m pytest tests -m gpu -q --maxfail=30 > $OUT/pytest_gpu_${TAG}_epg.log 2>&1
echo "pytest epg rc=$?"; tail -30 $OUT/pytest_gpu_${TAG}_epg.log | cut -c1-300
one() {  # name envs  (env vars from the caller)
  timeout 300 python bench.py --envs $2 --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>>$OUT/bench_$TAG.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('bench', '$1', $2, d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['launches_per_step'], d['scan_timeouts'])"
}
for rep in 1 2; do
  one epg16 65536
  LGK_TAIL_EPG=8 one epg8 65536
  LGK_TAIL_EPG=4 one epg4 65536
  LGK_TAIL_EPG=8 LGK_CHAIN=2w one epg8_wait 65536
  LGK_CHAIN=3 one chain3 65536
done
for v in ${VARIANTS:-}; do
  for e in 16 8; do LGK_TAIL_EPG=$e LGK_LIB_PATH=$PWD/build/liblgk_$v.so one ${v}_epg$e 65536; done
done
for e in 4096 16384; do
  one epg16 $e
  LGK_TAIL_EPG=8 one epg8 $e
done
tail -5 $OUT/bench_$TAG.err
