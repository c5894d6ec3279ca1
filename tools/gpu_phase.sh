#!/usr/bin/env bash
# GPU-box call: phase timings of the step kernel (profile build) for a few LGK_SCAN_CTAS, plus plain bench lines
set -u
mkdir -p gpurun_out
for s in ${SCAN_LIST:-6 8}; do
  echo "== LGK_SCAN_CTAS=$s"
  LGK_SCAN_CTAS=$s timeout 300 python tools/phase_times2.py 65536 2>&1 | tail -12
  LGK_SCAN_CTAS=$s timeout 300 python bench.py --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('bench', d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'])"
done
