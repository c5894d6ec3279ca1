#!/usr/bin/env bash
# kernel time of the fused step vs envs per persistent CTA. usage: gpurun -- 'bash tools/sweep_per_cta.sh "32 64 76 96"'
for p in $1; do
  LGK_ENVS_PER_CTA=$p python bench.py --no-cpu-baseline --steps 200 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('per_cta', $p, 'ms_per_step', d['ms_per_step'], 'kernel_ms', d['roofline']['kernel_ms'])"
done
