#!/usr/bin/env bash
# GPU-box call: scan group size sweep at 65 536 envs (LGK_SCAN_EPG), then the GPU tests on the default
set -u
OUT=gpurun_out; mkdir -p $OUT
for e in default 8 7 6 4 2 1; do
  if [ $e = default ]; then unset LGK_SCAN_EPG; else export LGK_SCAN_EPG=$e; fi
  timeout 300 python bench.py --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>>$OUT/bench_epg.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('epg', '$e', d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d.get('scan_timeouts'))"
done
unset LGK_SCAN_EPG
timeout 1500 python -m pytest tests -m gpu -q -x > $OUT/pytest_gpu_epg.log 2>&1
echo "pytest rc=$?"; tail -4 $OUT/pytest_gpu_epg.log | cut -c1-300
