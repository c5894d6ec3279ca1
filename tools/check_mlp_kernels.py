"""Go1 actuator MLP: the three kernels (fp32 FFMA, mma.sync 3xTF32, tcgen05 3xTF32) on the same inputs --
max |difference| of dVel against the FFMA kernel and a float64 torch evaluation, then median launch times.
usage: gpurun -- 'python tools/check_mlp_kernels.py [envs ...]'"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import helpers  # noqa: E402
from legged_gym_b200.utils import synthetic  # noqa: E402

KERNELS = ("ffma", "mma", "umma")


def f64_eval(env):
    w = env.actuator_weights.double().cpu()
    x = env.model_ins.double().cpu().reshape(-1, 30)
    o = 0
    layers = []
    for n_out, n_in in ((128, 30), (128, 128), (128, 128), (3, 128)):
        W = w[o:o + n_out * n_in].reshape(n_out, n_in); o += n_out * n_in
        b = w[o:o + n_out]; o += n_out
        layers.append((W, b))
    for i, (W, b) in enumerate(layers):
        x = x @ W.T + b
        if i < 3:
            x = torch.tanh(x)
    vel_std = env._act_norm.double().cpu()[36:48].reshape(4, 3)
    rows = x.shape[0]
    return (x.reshape(rows // 4, 4, 3) * vel_std).reshape(rows // 4, 12)


def main():
    sizes = [int(a) for a in sys.argv[1:]] or [1000, 16384, 65536]
    for N in sizes:
        env = helpers.make_env("go1_rough_actuator", N, device="cuda:0", seed=1)
        fr = synthetic.make_frames(env, 1, seed=3)[0]
        synthetic.load_frame(env, fr)
        act = fr["actions"].clone()
        out = {}
        for k in KERNELS:
            os.environ["LGK_MLP_KERNEL"] = k
            env.pos_err_buffs.zero_(); env.vel_buffs.zero_()
            out[k] = env.actuator_advance(act).clone()
            torch.cuda.synchronize()
        want = f64_eval(env).to("cuda:0")
        scale = float(want.abs().max())
        for k in KERNELS:
            print(f"envs {N:6d} {k:5s}: max|dVel - f64| {float((out[k].double() - want).abs().max()):.3e}   "
                  f"max|dVel - ffma| {float((out[k] - out['ffma']).abs().max()):.3e}   (max|dVel| {scale:.3f})", flush=True)
        for k in KERNELS:
            os.environ["LGK_MLP_KERNEL"] = k
            for _ in range(5):
                env.actuator_advance(act)
            torch.cuda.synchronize()
            ts = []
            for _ in range(30):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); env.actuator_advance(act); e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            ts.sort()
            print(f"envs {N:6d} {k:5s}: {ts[len(ts) // 2] * 1e3:8.1f} us per substep (history shift + MLP, median of 30)", flush=True)
        del env


if __name__ == "__main__":
    main()
