"""Development sweep on the GPU box: parity of every scenario against the oracle on torch-CPU and
torch-CUDA; prints failures instead of stopping at the first."""
import os
import sys
import traceback

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
from legged_gym_b200.utils import scenarios  # noqa: E402

if __name__ == "__main__":
    print(torch.cuda.get_device_name(0))
    envs = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
    for dev in ("cpu", "cuda:0"):
        for name in scenarios.SCENARIOS:
            try:
                n = helpers.run_parity(name, envs, 6, oracle_device=dev)
                print(f"PASS {name} vs oracle[{dev}] ({n} resets)")
            except Exception as e:   # noqa: BLE001
                print(f"FAIL {name} vs oracle[{dev}]: {str(e)[:400]}")
                if not isinstance(e, AssertionError):
                    traceback.print_exc()
