#!/usr/bin/env bash
# GPU-box call: the GPU tests without programmatic dependent launch (LGK_NO_PDL=1), then the default suite and bench lines
set -u
TAG=${1:-r03g}
OUT=gpurun_out; mkdir -p $OUT
LGK_NO_PDL=1 timeout 800 python -m pytest tests -x -q -m gpu > $OUT/${TAG}_pytest_gpu_nopdl.txt 2>&1; echo "pytest nopdl rc=$?"; tail -1 $OUT/${TAG}_pytest_gpu_nopdl.txt
timeout 800 python -m pytest tests -x -q -m gpu > $OUT/${TAG}_pytest_gpu.txt 2>&1; echo "pytest rc=$?"; tail -1 $OUT/${TAG}_pytest_gpu.txt
one() {
  timeout 300 python bench.py --envs $2 --steps 200 --warmup 20 --no-cpu-baseline --no-e2e 2>>$OUT/bench_$TAG.err | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('bench', '$1', $2, d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['launches_per_step'], d['scan_timeouts'])"
}
for rep in 1 2 3; do one default 65536; done
LGK_NO_PDL=1 one nopdl 65536
LGK_CHAIN=2w one wait 65536
LGK_CHAIN=3 one chain3 65536
one default 4096
one default 16384
