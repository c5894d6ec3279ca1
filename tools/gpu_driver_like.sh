#!/usr/bin/env bash
# GPU-box call (--gpus G): the round-end driver's own bench commands (--steps 20 --warmup 5), REPS times per N
set -u
G=${1:-2}; REPS=${2:-3}; TAG=${3:-r02s}
OUT=gpurun_out; mkdir -p $OUT
for n in 1 $G; do
  for r in $(seq 1 $REPS); do
    if [ $n = 1 ]; then CMD="python bench.py --gpus 1 --steps 20 --warmup 5"
    else CMD="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + r)) bench.py --gpus $n --steps 20 --warmup 5"; fi
    LGK_BENCH_WATCHDOG=240 timeout 300 $CMD > $OUT/${TAG}_n${n}_$r.json 2> $OUT/${TAG}_n${n}_$r.err
    echo "n=$n rep=$r rc=$?"
    grep '^{' $OUT/${TAG}_n${n}_$r.json | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print('  ', d['n_gpus'], d['ms_per_step'], '%.4g' % d['value'], 'e2e %.4g' % d['e2e']['value'], 'coll', d.get('collectives_in_timed_region'), 'clk', d['clocks'], 'frac', d['roofline']['frac'])"
  done
done
