"""Which LgkIndexMath mode reproduces torch's height-field indices bit-for-bit on this machine?
Runs the fused kernel in every (division, norm) mode against the oracle on torch-CPU and torch-CUDA
and prints the number of measured_heights mismatches.  GPU box only.

    python tools/calibrate_index_math.py [--envs 16384]
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import helpers  # noqa: E402
from legged_gym_b200.utils import synthetic  # noqa: E402
from oracle import legged_oracle as lo  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=16384)
    ap.add_argument("--frames", type=int, default=4)
    a = ap.parse_args()
    env = helpers.make_env("go1_rough_default", a.envs, device="cuda:0")
    frames = synthetic.make_frames(env, a.frames, seed=7)
    results = {}
    for dev in ("cpu", "cuda:0"):
        p = lo.OracleParams.from_env(env, device=dev)
        for fi, fr in enumerate(frames):
            s = lo.OracleState.from_env(env, p, device=dev)
            s.root_states.copy_(fr["root_states"].to(dev))
            want = lo.get_heights(p, s).cpu()
            for mode in range(2):
                synthetic.load_frame(env, fr)
                taps = synthetic.make_taps(env, seed=fi)
                env.use_rng_taps(taps, index_math=mode)
                env.episode_length_buf = torch.ones(a.envs, dtype=torch.long, device="cuda:0")
                env.step(fr["actions"].clone())
                torch.cuda.synchronize()
                bad = int((env.measured_heights.cpu() != want).sum())
                results[(dev, mode)] = results.get((dev, mode), 0) + bad
    total = a.envs * env.num_height_points * a.frames
    for dev in ("cpu", "cuda:0"):
        for mode in range(2):
            print(f"oracle on {dev:6s}  div={'mul_rcp' if mode & 1 else 'ieee_div':8s} "
                  f"(index_math={mode}): {results[(dev, mode)]} / {total} heights differ")


if __name__ == "__main__":
    main()
