"""Runs the reference's own code (through the shim) and the restated oracle side by side on the same
synthetic frames and taps, and requires BIT-IDENTICAL buffers on the same torch device.
TEST INFRASTRUCTURE; only runs where /root/reference exists.

    python -m oracle.validate_against_reference [--envs 256] [--steps 12]
"""
from __future__ import annotations

import argparse
import sys

import torch

from legged_gym_b200.utils import scenarios, synthetic
from oracle import legged_oracle as lo
from oracle.reference_loader import TapHarness, make_reference_env

CHECK = ("obs_buf", "rew_buf", "reset_buf", "time_out_buf", "episode_length_buf", "commands", "root_states",
         "dof_state", "last_actions", "last_dof_vel", "last_root_vel", "feet_air_time", "base_lin_vel",
         "base_ang_vel", "projected_gravity", "env_origins")


def compare_step(env, s, p, tag):
    bad = []
    for name in CHECK:
        a = getattr(env, name)
        b = getattr(s, name)
        if not torch.equal(a.reshape(-1), b.reshape(-1).to(a.dtype)):
            bad.append(name)
    if hasattr(env, "terrain_levels") and not torch.equal(env.terrain_levels, s.terrain_levels):
        bad.append("terrain_levels")
    for k in env.episode_sums:
        if not torch.equal(env.episode_sums[k], s.episode_sums[k]):
            bad.append("episode_sums." + k)
    if torch.is_tensor(env.measured_heights) and not torch.equal(env.measured_heights, s.measured_heights):
        bad.append("measured_heights")
    ea, eb = env.extras.get("episode", {}), s.extras.get("episode", {})
    for k in ea:
        va, vb = torch.as_tensor(ea[k]), torch.as_tensor(eb[k])
        if not torch.equal(va.float(), vb.float()):
            bad.append("extras." + k)
    for k, v in env.command_ranges.items():   # the command curriculum widens lin_vel_x in place (:465-474)
        if [float(x) for x in v] != [float(x) for x in p.command_ranges[k]]:
            bad.append("command_ranges." + k)
    if bad:
        raise AssertionError(f"{tag}: oracle != reference in {bad}")


def run_scenario(name, num_envs=256, steps=12, seed=0, verbose=True):
    task, mutate = scenarios.SCENARIOS[name]
    env = make_reference_env(task, mutate, num_envs=num_envs, seed=seed)
    p = lo.OracleParams.from_env(env)
    frames = synthetic.make_frames(env, 4, seed=1234 + seed)
    # start from a mid-episode state so resampling / time-outs / pushes fire within a few steps
    g = torch.Generator().manual_seed(seed)
    env.episode_length_buf[:] = torch.randint(0, int(env.max_episode_length) + 2, (num_envs,), generator=g)
    env.episode_length_buf[: num_envs // 8] = 499          # callback resample on the first step
    env.common_step_counter = scenarios.fixture_step_counter(env)   # a push (and the command-curriculum gate) on the third step
    env.commands[:] = torch.rand(num_envs, 4, generator=g) * 2 - 1
    env.feet_air_time[:] = torch.rand(num_envs, env.feet_air_time.shape[1], generator=g) * 0.6
    env.last_actions[:] = torch.randn(num_envs, env.num_actions, generator=g)
    env.last_dof_vel[:] = torch.randn(num_envs, env.num_dof, generator=g)
    for k in env.episode_sums:
        env.episode_sums[k][:] = torch.randn(num_envs, generator=g)
    scenarios.fixture_bias_tracking_sums(env, env.episode_sums)
    synthetic.load_frame(env, frames[0])
    s = lo.OracleState.from_env(env, p)
    harness = TapHarness(env)
    n_resets = 0
    for it in range(steps):
        fr = frames[it % len(frames)]
        taps = synthetic.make_taps(env, seed=100 * seed + it)
        # reference: step() with the frame written in place of simulate()
        synthetic.load_frame(env, fr)
        with harness.active(taps):
            env.step(fr["actions"].clone())
        # oracle
        s.root_states.copy_(fr["root_states"]); s.dof_state.copy_(fr["dof_state"])
        s.contact_forces.copy_(fr["contact_forces"].view_as(s.contact_forces)); s.torques.copy_(fr["torques"])
        target = lo.step(p, s, fr["actions"].clone(), taps)
        compare_step(env, s, p, f"{name} step {it}")
        assert torch.equal(env.target_poses, target)
        n_resets += int(env.reset_buf.sum())
    if verbose:
        print(f"  {name}: {steps} steps x {num_envs} envs bit-identical ({n_resets} resets, "
              f"rewards {list(env.reward_names)} + {'termination' if 'termination' in env.reward_scales else '-'})")
    return n_resets


def run_compute_torques(num_envs=96, seed=0, verbose=True):
    """LeggedRobot._compute_torques (legged_robot.py:370-392) of the reference's base class for the three control
    types against lo.compute_torques, bit-identical; an unknown type raises NameError in both."""
    from legged_gym.envs.base.legged_robot import LeggedRobot
    env = make_reference_env("go1", scenarios.go1_flat, num_envs=num_envs, seed=seed)
    frames = synthetic.make_frames(env, 3, seed=4321 + seed)
    g = torch.Generator().manual_seed(seed)
    for it, ct in enumerate(("P", "V", "T")):
        synthetic.load_frame(env, frames[it])
        env.cfg.control.control_type = ct
        env.last_dof_vel[:] = torch.randn(num_envs, env.num_dof, generator=g)
        p = lo.OracleParams.from_env(env)
        s = lo.OracleState.from_env(env, p)
        assert p.control_type == ct
        a = torch.randn(num_envs, env.num_actions, generator=g) * 3.0     # some torques hit the limits
        want = LeggedRobot._compute_torques(env, a)                       # the base class's, not Go1's override
        got = lo.compute_torques(p, s, a)
        assert torch.equal(want, got), f"control_type {ct}: oracle != reference"
        assert ct == "T" or bool((got.abs() == p.torque_limits).any())
    env.cfg.control.control_type = "X"
    p = lo.OracleParams.from_env(env)
    for fn in (lambda: LeggedRobot._compute_torques(env, a), lambda: lo.compute_torques(p, lo.OracleState.from_env(env, p), a)):
        try:
            fn()
        except NameError:
            pass
        else:
            raise AssertionError("unknown control type must raise NameError (:391)")
    if verbose:
        print(f"  _compute_torques P / V / T: {num_envs} envs bit-identical, unknown type -> NameError")
    return True


def run_sea_lstm(num_envs=64, steps=20, seed=0, verbose=True):
    """ANYmal actuator net: the reference's Anymal._compute_torques / reset_idx with the SHIPPED TorchScript
    checkpoint (resources/actuator_nets/anydrive_v3_lstm.pt, aten::lstm) against SeaLstmOracle's written-out
    cell.  aten::lstm's CPU kernel orders the gate sums differently, so this is not bit-identical: each call
    starts from the reference's state and must agree within a few ulp (state) / 1e-6 of the output-layer
    operand magnitude (torques); the free-running drift is reported."""
    env = make_reference_env("anymal_c_flat", None, num_envs=num_envs, seed=seed)
    assert env.cfg.control.use_actuator_network
    sd = dict(env.actuator_network.state_dict())
    p = lo.OracleParams.from_env(env)
    sea = lo.SeaLstmOracle(num_envs, env.num_actions, sd)
    free = lo.SeaLstmOracle(num_envs, env.num_actions, sd)
    atol = 1e-6 * float(sd["out_scale"].abs().max() * (sd["linear.bias"].abs().sum() + sd["linear.weight"].abs().sum()))
    g = torch.Generator().manual_seed(seed)
    frames = synthetic.make_frames(env, 4, seed=77 + seed)
    worst_state = worst_tq = drift = 0.0
    for it in range(steps):
        synthetic.load_frame(env, frames[it % 4])
        a = torch.randn(num_envs, env.num_actions, generator=g)
        sea.sea_hidden_state.copy_(env.sea_hidden_state); sea.sea_cell_state.copy_(env.sea_cell_state)
        want = env._compute_torques(a)
        got = sea.compute_torques(a, p.action_scale, p.default_dof_pos, env.dof_pos, env.dof_vel)
        fr = free.compute_torques(a, p.action_scale, p.default_dof_pos, env.dof_pos, env.dof_vel)
        assert torch.equal(sea.sea_input, env.sea_input)
        worst_state = max(worst_state, float((sea.sea_hidden_state - env.sea_hidden_state).abs().max()),
                          float((sea.sea_cell_state - env.sea_cell_state).abs().max()))
        worst_tq = max(worst_tq, float(((got - want).abs() - 1e-5 * want.abs()).max()))
        drift = max(drift, float((free.sea_hidden_state - env.sea_hidden_state).abs().max()))
        if it == steps // 2:                    # reset_idx override zeroes exactly the listed envs (:56-60)
            ids = torch.tensor([0, 3, num_envs - 1])
            env.sea_hidden_state_per_env[:, ids] = 0.; env.sea_cell_state_per_env[:, ids] = 0.
            free.reset(ids)
    assert worst_state < 1e-6, worst_state
    assert worst_tq < atol, (worst_tq, atol)
    assert drift < 1e-4, drift
    if verbose:
        print(f"  anymal LSTM (shipped checkpoint): {steps} calls x {num_envs * env.num_actions} actuators, per-call "
              f"state error {worst_state:.1e}, torque error beyond 1e-5 rel {max(worst_tq, 0):.1e} (floor {atol:.1e}), "
              f"free-running drift {drift:.1e}")
    return worst_state


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=256)
    ap.add_argument("--steps", type=int, default=12)
    a = ap.parse_args(argv)
    for name in scenarios.SCENARIOS:
        run_scenario(name, a.envs, a.steps)
    run_compute_torques()
    run_sea_lstm()
    print("oracle == reference on every scenario")


if __name__ == "__main__":
    sys.exit(main())
