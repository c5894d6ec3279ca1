"""Times the ANYmal LSTM actuator substep (Anymal._compute_torques, anymal.py:71-81)
and, beside it, the reference's way of computing the same thing on the same GPU: a torch.nn.LSTM(2, 8, 2
layers) + Linear(8, 1) module with the same weights called on [N*12, 1, 2] inputs (cuDNN RNN path).
Prints one JSON line per env count.  usage: gpurun -- 'python tools/bench_lstm.py [envs ...]'"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import helpers  # noqa: E402
from legged_gym_b200.utils import synthetic  # noqa: E402

BYTES_PER_ACTUATOR = 4 + 8 + 4 * 32 + 4 * 32 + 4 + 8     # action, dof_state, state in, state out, torque, sea_input
FLOP_PER_ACTUATOR = 2 * (32 * 2 + 3 * 32 * 8 + 8)


def timed(fn, reps=100, flush=None):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        if flush is not None:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    sizes = [int(a) for a in sys.argv[1:]] or [4096, 65536]
    peak = 6442.9
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:   # noqa: BLE001
        pass
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
    for N in sizes:
        env = helpers.make_env("anymal_c_flat", N, device="cuda:0", seed=1)
        fr = synthetic.make_frames(env, 1, seed=3)[0]
        synthetic.load_frame(env, fr)
        act = fr["actions"].clone()
        ms = timed(lambda: env._compute_torques(act), flush=flush)
        M = N * env.num_actions

        class LSTMsea(torch.nn.Module):
            def __init__(self):
                super().__init__()
                self.lstm = torch.nn.LSTM(2, 8, num_layers=2, batch_first=True)
                self.linear = torch.nn.Linear(8, 1)
                self.register_buffer("in_scale", torch.ones(1, 1, 2))
                self.register_buffer("out_scale", torch.ones(1))

            def forward(self, x, hc0):
                x1, hcn = self.lstm(torch.mul(x, self.in_scale), hc0)
                return torch.mul(self.out_scale, torch.squeeze(self.linear(x1))), hcn
        net = LSTMsea()
        net.load_state_dict(env.actuator_state_dict)
        net = net.to("cuda:0")
        h = torch.zeros(2, M, 8, device="cuda:0"); c = torch.zeros(2, M, 8, device="cuda:0")
        sea_input = torch.zeros(M, 1, 2, device="cuda:0")

        def torch_way():
            with torch.inference_mode():
                sea_input[:, 0, 0] = (act * env.cfg.control.action_scale + env.default_dof_pos - env.dof_pos).flatten()
                sea_input[:, 0, 1] = env.dof_vel.flatten()
                t, (h[:], c[:]) = net(sea_input, (h, c))
            return t
        ms_t = timed(torch_way, flush=flush)
        gbs = BYTES_PER_ACTUATOR * M / (ms * 1e-3) / 1e9
        print(json.dumps({
            "kernel": "lgk_lstm_kernel", "envs": N, "actuators": M, "us_per_substep": round(ms * 1e3, 2),
            "env_substeps_per_s": round(N / (ms * 1e-3), 1), "algorithmic_bytes_per_actuator": BYTES_PER_ACTUATOR,
            "achieved_gbs": round(gbs, 1), "hbm_peak_gbs": peak, "frac_hbm": round(gbs / peak, 4),
            "tflops_fp32": round(FLOP_PER_ACTUATOR * M / (ms * 1e-3) / 1e12, 2),
            "torch_cuda_same_module_us": round(ms_t * 1e3, 2), "speedup_vs_torch_cuda": round(ms_t / ms, 2),
            "timing": "median of 100 single launches, L2 flushed (256 MB memset) before each"}))


if __name__ == "__main__":
    main()
