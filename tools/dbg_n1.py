import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, helpers
from legged_gym_b200 import _lgk
from legged_gym_b200.utils import synthetic
from oracle import legged_oracle as lo
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1
seed = 2
env = helpers.make_env("go1_rough_full", N, device="cuda:0", seed=seed)
helpers.randomise_env_state(env, seed)
frames = synthetic.make_frames(env, 4, seed=1234 + seed)
synthetic.load_frame(env, frames[0])
p = lo.OracleParams.from_env(env, device="cpu")
s = lo.OracleState.from_env(env, p, device="cpu")
for it in range(4):
    fr = frames[it % 4]
    taps = synthetic.make_taps(env, seed=100 * seed + it)
    synthetic.load_frame(env, fr)
    env.use_rng_taps(taps, index_math=0)
    print("step", it, "fat before", env.feet_air_time[0].tolist(), s.feet_air_time[0].tolist())
    print("  feet fz", env.contact_forces[0, env.feet_indices, 2].tolist(), "cmd", env.commands[0].tolist())
    env.step(fr["actions"].clone())
    torch.cuda.synchronize()
    od = {k: v.cpu() for k, v in fr.items()}
    s.root_states.copy_(od["root_states"]); s.dof_state.copy_(od["dof_state"])
    s.contact_forces.copy_(od["contact_forces"].view_as(s.contact_forces)); s.torques.copy_(od["torques"])
    lo.step(p, s, od["actions"].clone(), {k: v.cpu() for k, v in taps.items()})
    print("  fat after ", env.feet_air_time[0].tolist(), s.feet_air_time[0].tolist(), "reset", env.reset_buf[0].item(), s.reset_buf[0].item())
