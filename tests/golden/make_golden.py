"""Generates tests/golden/<scenario>.npz by running the REFERENCE's own, unmodified code
(/root/reference, imported over the fake-Isaac-Gym shim in oracle/refshim) on seeded synthetic
simulator frames and RNG taps, and recording every buffer it produces.

Only runs in the build container (the GPU box has no /root/reference); the fixtures travel.

    python tests/golden/make_golden.py [--envs 64] [--steps 6]

Fixture layout (numpy arrays; S = steps, N = envs):
  meta_*            scenario name, N, S, seed, crc32 of the terrain height field / origins
  init_*            env state before the first step (episode_length_buf, commands, feet_air_time,
                    last_actions, last_dof_vel, episode_sums [K,N], terrain_levels, env_origins,
                    common_step_counter)
  in_<name>         [S, ...] simulator frames + actions fed to step s
  tap_<name>        [S, ...] uniform taps consumed by step s
  out_<name>        [S, ...] the reference's buffers after step s
  out_reset_ids / out_resample_ids   [S, N] ascending ids padded with -1
  out_extras        [S, K+1] extras["episode"] values in episode_sums key order (+ terrain_level), NaN
                    where the reference did not rewrite them
  out_command_ranges [S, 4, 2] command_ranges (lin_vel_x, lin_vel_y, ang_vel_yaw, heading) after step s: the command
                    curriculum (legged_robot.py:465-474) widens lin_vel_x on its gated step
  out_dvel          [S, decimation, N, 12] Go1 actuator-net output per substep (actuator scenario only)
  in_sea_dof_state / out_sea_torques / out_sea_hidden / out_sea_cell / out_sea_input
                    ANYmal scenario only: after step s the reference's Anymal._compute_torques(actions) is
                    called `decimation` times on the dof states in_sea_dof_state[s, k]; torques per call
                    [S, dec, N*12], LSTM state / sea_input after the last call of the step [S, 2, N*12, 8] /
                    [S, N*12, 1, 2] (the resets of step s have zeroed the state of the reset envs through
                    the reference's reset_idx override before the calls)
"""
from __future__ import annotations

import argparse
import os
import sys
import zlib

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from legged_gym_b200.utils import scenarios, synthetic  # noqa: E402
from oracle.reference_loader import TapHarness, make_reference_env  # noqa: E402

OUT_BUFFERS = ("obs_buf", "rew_buf", "reset_buf", "time_out_buf", "episode_length_buf", "commands", "root_states",
               "dof_state", "last_actions", "last_dof_vel", "last_root_vel", "feet_air_time", "base_lin_vel",
               "base_ang_vel", "projected_gravity", "env_origins", "target_poses")
IN_BUFFERS = ("root_states", "dof_state", "contact_forces", "torques", "actions")


def crc(t) -> int:
    return zlib.crc32(np.ascontiguousarray(t.detach().cpu().numpy()).tobytes()) & 0xFFFFFFFF


def np_(t):
    return t.detach().cpu().numpy().copy()


def randomise(env, seed):
    """Mid-episode state so resampling (step 0), time-outs (steps 0 and 2) and a push (step 2) fire."""
    g = torch.Generator().manual_seed(seed)
    N = env.num_envs
    ep = torch.randint(0, int(env.max_episode_length) + 2, (N,), generator=g)
    ep[: max(N // 8, 1)] = int(env.cfg.commands.resampling_time / env.dt) - 1
    ep[N // 8: N // 8 + 3] = int(env.max_episode_length)            # time-outs on the first step
    ep[N // 8 + 3: N // 8 + 5] = int(env.max_episode_length) - 2    # ... and on the third
    env.episode_length_buf[:] = ep
    env.common_step_counter = scenarios.fixture_step_counter(env)
    env.commands[:] = torch.rand(N, 4, generator=g) * 2 - 1
    env.feet_air_time[:] = torch.rand(N, env.feet_air_time.shape[1], generator=g) * 0.6
    env.last_actions[:] = torch.randn(N, env.num_actions, generator=g)
    env.last_dof_vel[:] = torch.randn(N, env.num_dof, generator=g)
    for k in env.episode_sums:
        env.episode_sums[k][:] = torch.randn(N, generator=g)
    scenarios.fixture_bias_tracking_sums(env, env.episode_sums)


def generate(name, num_envs, steps, seed=0):
    task, mutate = scenarios.SCENARIOS[name]
    env = make_reference_env(task, mutate, num_envs=num_envs, seed=seed)
    actuator = task == "go1" and bool(env.cfg.control.use_actuator_network)
    dvel_log = []
    if actuator:
        # swap the shipped checkpoint for seeded random weights of the same architecture: the checkpoint
        # does not travel to the GPU box, the architecture and the arithmetic are what is under test
        from legged_gym_b200.envs.go1.go1 import random_mlp_state_dict
        sd = random_mlp_state_dict(0)
        net = torch.nn.Sequential(torch.nn.Linear(30, 128), torch.nn.Tanh(), torch.nn.Linear(128, 128), torch.nn.Tanh(),
                                  torch.nn.Linear(128, 128), torch.nn.Tanh(), torch.nn.Linear(128, 3))
        net.load_state_dict({k.replace("architecture.", ""): v for k, v in sd.items()})
        env.actuator_network.core_model = net
        orig = env.actuator_advance

        def logged(actions):
            out = orig(actions)
            dvel_log.append(np_(out))
            return out
        env.actuator_advance = logged
    anymal = task.startswith("anymal") and bool(env.cfg.control.use_actuator_network)
    sea = {k: [] for k in ("in_sea_dof_state", "out_sea_torques", "out_sea_hidden", "out_sea_cell", "out_sea_input")}
    if anymal:
        # the reference loaded its checkpoint (anymal.py:51-54); swap in seeded random weights of the same
        # architecture (LSTMsea: x*in_scale -> LSTM(2,8,2 layers, batch_first) -> Linear(8,1) -> *out_scale)
        from legged_gym_b200.envs.anymal_c.anymal import random_lstm_state_dict

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
        assert sorted(env.actuator_network.state_dict().keys()) == sorted(random_lstm_state_dict(0).keys())
        net = LSTMsea()
        net.load_state_dict(random_lstm_state_dict(0))
        env.actuator_network = net
    randomise(env, seed)
    frames = synthetic.make_frames(env, steps, seed=1234 + seed)
    synthetic.load_frame(env, frames[0])
    keys = list(env.episode_sums.keys())
    d = {"meta_scenario": np.array(name), "meta_envs": np.array(num_envs), "meta_steps": np.array(steps),
         "meta_seed": np.array(seed), "meta_sum_keys": np.array(keys),
         "meta_reward_names": np.array(list(env.reward_names)),
         "meta_terrain_crc": np.array(crc(env.height_samples) if getattr(env, "height_samples", None) is not None else 0),
         "meta_origins_crc": np.array(crc(env.terrain_origins) if hasattr(env, "terrain_origins") else 0)}
    for k in ("episode_length_buf", "commands", "feet_air_time", "last_actions", "last_dof_vel", "env_origins"):
        d["init_" + k] = np_(getattr(env, k))
    d["init_episode_sums"] = np.stack([np_(env.episode_sums[k]) for k in keys])
    d["init_terrain_levels"] = np_(env.terrain_levels) if hasattr(env, "terrain_levels") else np.zeros(0, np.int64)
    d["init_terrain_types"] = np_(env.terrain_types) if hasattr(env, "terrain_types") else np.zeros(0, np.int64)
    d["init_common_step_counter"] = np.array(env.common_step_counter)
    harness = TapHarness(env)
    ins = {k: [] for k in IN_BUFFERS}
    taps_log = {}
    outs = {k: [] for k in OUT_BUFFERS + ("episode_sums", "terrain_levels", "measured_heights", "reset_ids",
                                          "resample_ids", "extras", "command_ranges")}
    interval = int(env.cfg.commands.resampling_time / env.dt)
    for it in range(steps):
        fr = frames[it]
        taps = synthetic.make_taps(env, seed=100 * seed + it)
        for k in IN_BUFFERS:
            ins[k].append(np_(fr[k]))
        for k, v in taps.items():
            taps_log.setdefault(k, []).append(np_(v))
        resample = ((env.episode_length_buf + 1) % interval == 0).nonzero(as_tuple=False).flatten()   # :118,:342
        synthetic.load_frame(env, fr)
        env.extras.pop("episode", None)
        with harness.active(taps):
            env.step(fr["actions"].clone())
        for k in OUT_BUFFERS:
            outs[k].append(np_(getattr(env, k)))
        outs["episode_sums"].append(np.stack([np_(env.episode_sums[k]) for k in keys]))
        outs["terrain_levels"].append(np_(env.terrain_levels) if hasattr(env, "terrain_levels") else np.zeros(0, np.int64))
        mh = env.measured_heights
        outs["measured_heights"].append(np_(mh) if torch.is_tensor(mh) else np.zeros((num_envs, 0), np.float32))
        pad = lambda ids: np.concatenate([np_(ids).astype(np.int64), -np.ones(num_envs - len(ids), np.int64)])
        outs["reset_ids"].append(pad(env.reset_buf.nonzero(as_tuple=False).flatten()))
        outs["resample_ids"].append(pad(resample))
        if anymal:
            g_sea = torch.Generator().manual_seed(7000 + 10 * seed + it)
            for k in range(env.cfg.control.decimation):
                ds = fr["dof_state"].clone().view(num_envs, env.num_dof, 2)
                ds[..., 0] += 0.05 * torch.randn(num_envs, env.num_dof, generator=g_sea)
                ds[..., 1] += 0.5 * torch.randn(num_envs, env.num_dof, generator=g_sea)
                env.dof_state.copy_(ds.view_as(env.dof_state))
                tq = env._compute_torques(env.actions)
                sea["in_sea_dof_state"].append(np_(ds))
                sea["out_sea_torques"].append(np_(tq))
            sea["out_sea_hidden"].append(np_(env.sea_hidden_state))
            sea["out_sea_cell"].append(np_(env.sea_cell_state))
            sea["out_sea_input"].append(np_(env.sea_input))
        ep = env.extras.get("episode", {})
        row = [float(ep["rew_" + k]) if ("rew_" + k) in ep else float("nan") for k in keys]
        row.append(float(ep["terrain_level"]) if "terrain_level" in ep else float("nan"))
        outs["extras"].append(np.array(row, np.float32))
        outs["command_ranges"].append(np.array([env.command_ranges[k] for k in ("lin_vel_x", "lin_vel_y", "ang_vel_yaw", "heading")],
                                               np.float64))
    for k, v in ins.items():
        d["in_" + k] = np.stack(v)
    for k, v in taps_log.items():
        d["tap_" + k] = np.stack(v)
    for k, v in outs.items():
        d["out_" + k] = np.stack(v)
    if actuator:
        dec = env.cfg.control.decimation
        d["out_dvel"] = np.stack(dvel_log).reshape(steps, dec, num_envs, 12)
        d["out_model_ins"] = np_(env.model_ins)
    if anymal:
        dec = env.cfg.control.decimation
        for k, v in sea.items():
            a = np.stack(v)
            d[k] = a.reshape((steps, dec) + a.shape[1:]) if a.shape[0] == steps * dec else a
    n_reset = int((d["out_reset_ids"] >= 0).sum())
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **d)
    print(f"  {name}: {steps} steps x {num_envs} envs, {n_reset} resets, {os.path.getsize(path) / 1e3:.0f} kB -> {path}")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=64)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--only", default=None, help="comma-separated scenario names (default: all)")
    a = ap.parse_args()
    for name in scenarios.SCENARIOS:
        if a.only is None or name in a.only.split(","):
            generate(name, a.envs, a.steps)


if __name__ == "__main__":
    main()
