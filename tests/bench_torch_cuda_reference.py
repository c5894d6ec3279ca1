"""Not a test: times the reference's op-by-op torch implementation of the path ON THE GPU (the oracle
restatement, validated bit-identical to the reference on CPU torch, run with device='cuda:0') -- the
"same box" competitor SURVEY.md 8(d) asks to report next to the fused kernels.  Lives under tests/ because
only test infrastructure may import oracle/.

    gpurun -- 'python tests/bench_torch_cuda_reference.py [envs ...]'   -> one JSON line per env count
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import helpers  # noqa: E402
from legged_gym_b200.utils import synthetic  # noqa: E402
from oracle import legged_oracle as lo  # noqa: E402


def main():
    sizes = [int(a) for a in sys.argv[1:]] or [4096, 65536]
    dev = "cuda:0"
    for n in sizes:
        env = helpers.make_env("go1_rough_full", n, device="cpu", seed=1)
        g = torch.Generator().manual_seed(99)
        env.episode_length_buf = torch.randint(0, int(env.max_episode_length), (n,), generator=g)
        p = lo.OracleParams.from_env(env, device=dev)
        s = lo.OracleState.from_env(env, p, device=dev)
        frames = [{k: v.to(dev) for k, v in fr.items()} for fr in synthetic.make_frames(env, 4, seed=3234)]
        taps = {k: v.to(dev) for k, v in synthetic.make_taps(env, seed=5).items()}
        steps, warm = 30, 5
        ts = []
        for i in range(steps + warm):
            fr = frames[i % 4]
            s.root_states.copy_(fr["root_states"]); s.dof_state.copy_(fr["dof_state"])
            s.contact_forces.copy_(fr["contact_forces"].view_as(s.contact_forces)); s.torques.copy_(fr["torques"])
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            taps["noise"] = torch.rand(n, env.num_obs, device=dev)      # the reference draws its noise every step (:231)
            lo.step(p, s, fr["actions"], taps)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts = sorted(ts[warm:])
        ms = ts[len(ts) // 2]
        print(json.dumps({"impl": "reference op-by-op on CUDA torch (oracle restatement, cuda:0)", "envs": n,
                          "ms_per_step": round(ms, 3), "env_steps_per_s": round(n / (ms * 1e-3), 1),
                          "torch": torch.__version__, "note": "median of 30 steps, includes the reference's host syncs "
                          "(nonzero / len on the reset and resample id lists)"}))


if __name__ == "__main__":
    main()
