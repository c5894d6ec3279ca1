#!/usr/bin/env bash
# GPU-box call: one full bench line per BASELINE.json config (cpu_baseline at the config's own size, torch-CUDA
# competitor, e2e breakdown), the headline line, and the reference arm.
set -u
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
run() {  # name scenario envs
  timeout 900 python bench.py --scenario $2 --envs $3 --steps 200 --warmup 20 --cpu-budget 10 > $OUT/${TAG}_bench_$1.json 2> $OUT/${TAG}_bench_$1.err
  echo "== $1 rc=$?"; python - <<PY
import json
try:
    d = json.loads(open("$OUT/${TAG}_bench_$1.json").read().strip().splitlines()[-1])
    r = d["roofline"]; c = d.get("cpu_baseline") or {}; t = d.get("torch_cuda") or {}
    print(" ms/step", d["ms_per_step"], " value %.3e" % d["value"], " e2e %.3e" % d["e2e"]["value"], " roofline", r.get("bound"), r.get("frac"),
          " cpu", c.get("kind"), c.get("value"), " torch_cuda", t.get("ms_per_step"), (t.get("kind") or "")[:40], t.get("reference_on_cuda_failed"))
    print(" e2e breakdown", d["e2e"].get("breakdown"))
except Exception as e:
    print(" parse failed", e); print(open("$OUT/${TAG}_bench_$1.err").read()[-1500:])
PY
}
run config1_go1_flat_1800 go1_flat 1800
run config2_go1_rough_4096 go1_rough_full 4096
run config3_go1_actuator_16384 go1_rough_actuator 16384
run config4_cassie_8192 cassie_flat 8192
run headline_65536 go1_rough_full 65536
timeout 600 python bench.py --impl reference --steps 20 --warmup 3 > $OUT/${TAG}_bench_reference_arm.json 2> $OUT/${TAG}_bench_reference_arm.err
echo "== reference arm rc=$?"; tail -c 700 $OUT/${TAG}_bench_reference_arm.json
