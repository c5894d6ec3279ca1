"""Times the Go1 actuator-net substep (history shift + 30-128-128-128-3 tanh MLP per leg)
on config 3's shape (16384 envs) and reports it against the fp32 FFMA peak of the device.
usage: gpurun -- 'python tools/bench_actuator.py [envs]'"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, helpers
from legged_gym_b200.utils import synthetic
N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
env = helpers.make_env("go1_rough_actuator", N, device="cuda:0", seed=1)
fr = synthetic.make_frames(env, 1, seed=3)[0]
synthetic.load_frame(env, fr)
act = fr["actions"].clone()
for _ in range(5):
    env.actuator_advance(act)
torch.cuda.synchronize()
reps = 50
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    env.actuator_advance(act)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
flop = 2 * (30 * 128 + 128 * 128 + 128 * 128 + 128 * 3) * 4 * N
p = torch.cuda.get_device_properties(0)
peak = p.multi_processor_count * 128 * 2 * 1.965e9
import os as _os
_k = "ffma" if _os.environ.get("LGK_MLP_FFMA") == "1" else _os.environ.get("LGK_MLP_KERNEL", "umma")
print("kernel:", {"f": "fp32 FFMA", "m": "3xTF32 tensor pipe (mma.sync m16n8k8)",
                  "u": "3xTF32 tcgen05.mma kind::tf32, activations + accumulator in tensor memory"}[_k[0]])
print(f"actuator substep, {N} envs: {ms * 1e3:.1f} us  {flop / ms / 1e9:.2f} TFLOP/s fp32 = {100 * flop / (ms * 1e-3) / peak:.1f}% of the FFMA peak "
      f"({peak / 1e12:.1f} TFLOP/s at 1965 MHz)  -> {N / ms * 1e3:.3e} env-substeps/s")

# roofline of the MLP kernel on the tensor pipe: every product runs as three TF32 passes (layer 0's K = 30 padded to 32);
# the peak is half the measured dense bf16 rate of MEASURED_PEAKS.json (TF32 runs at half the bf16 rate; the in-kernel
# micro-benchmark of DESIGN.md section 4 -- 64 cycles per 128x128x8 MMA -- is the same figure at the boost clock)
import json as _json
_tf32 = 3 * 2 * (32 * 128 + 128 * 128 + 128 * 128) * 4 * N
try:
    _pk = _json.load(open(_os.path.join(ROOT, "MEASURED_PEAKS.json")))
    _peak, _src = _pk["bf16_tflops_sustained"] / 2, "MEASURED_PEAKS.json bf16_tflops_sustained / 2"
except Exception:   # noqa: BLE001
    _peak, _src = 4096 * p.multi_processor_count * 1.965e9 / 1e12, "64 cycles per 128x128x8 TF32 MMA x SMs x 1965 MHz"
_ach = _tf32 / (ms * 1e-3) / 1e12
print(_json.dumps({"kernel": _k, "envs": N, "us_per_substep": round(ms * 1e3, 1), "timed": "history kernel + MLP kernel, back to back",
                   "roofline": {"bound": "tensor", "achieved": round(_ach, 1), "peak": round(_peak, 1), "unit": "TFLOP/s",
                                "frac": round(_ach / _peak, 3), "peak_source": _src,
                                "flops": "3 TF32 passes x 2 x (32x128 + 2 x 128x128) per leg-row; the 128->3 layer is on the FMA pipe"}}))
