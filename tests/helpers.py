"""Shared test helpers: env construction and the kernel-vs-oracle comparison."""
from __future__ import annotations

import copy

import numpy as np
import torch

import legged_gym_b200.envs  # noqa: F401  (registers tasks)
from legged_gym_b200.sim import gymapi, gymutil
from legged_gym_b200.utils import scenarios, synthetic
from legged_gym_b200.utils.helpers import class_to_dict
from legged_gym_b200.utils.task_registry import task_registry

# north-star tolerance for floating point outputs: |a-b| <= 1e-5 * max(|a|,|b|) + 1e-6
RTOL, ATOL = 1e-5, 1e-6


def make_env(scenario: str, num_envs: int, device: str = "cpu", seed: int = 0, extra=None, task_class=None):
    task, mutate = scenarios.SCENARIOS[scenario] if scenario in scenarios.SCENARIOS else scenarios.SHIPPED_TASKS[scenario]
    cfg = copy.deepcopy(task_registry.env_cfgs[task])
    cfg.env.num_envs = num_envs
    cfg.seed = seed
    if task in ("go1", "aliengo", "anymal_c_flat", "anymal_c_rough", "anymal_b"):
        cfg.control.actuator_net_random_seed = 0
    mutate(cfg)
    if extra is not None:
        extra(cfg)
    sp = gymapi.SimParams()
    gymutil.parse_sim_config(class_to_dict(cfg.sim), sp)
    sp.use_gpu_pipeline = device.startswith("cuda")
    np.random.seed(seed)
    torch.manual_seed(seed)
    cls = task_class or task_registry.get_task_class(task)
    return cls(cfg=cfg, sim_params=sp, physics_engine=gymapi.SIM_PHYSX, sim_device=device, headless=True)


def randomise_env_state(env, seed: int):
    """Mid-episode state so that command resampling, time-outs and a push fire within a few steps."""
    g = torch.Generator().manual_seed(seed)
    N = env.num_envs
    dev = env.device
    ep = torch.randint(0, int(env.max_episode_length) + 2, (N,), generator=g)
    ep[: max(N // 8, 1)] = int(env.cfg.commands.resampling_time / env.dt) - 1
    env.episode_length_buf = ep.to(dev)
    env.common_step_counter = scenarios.fixture_step_counter(env)
    env.commands[:] = (torch.rand(N, 4, generator=g) * 2 - 1).to(dev)
    env.feet_air_time[:] = (torch.rand(N, env.feet_air_time.shape[1], generator=g) * 0.6).to(dev)
    env.last_actions[:] = torch.randn(N, env.num_actions, generator=g).to(dev)
    env.last_dof_vel[:] = torch.randn(N, env.num_dof, generator=g).to(dev)
    for k in env.episode_sums:
        env.episode_sums[k][:] = torch.randn(N, generator=g).to(dev)
    scenarios.fixture_bias_tracking_sums(env, env.episode_sums)


def assert_close(name, got, want, exact=False):
    got = got.detach().cpu()
    want = want.detach().cpu()
    assert got.shape == want.shape, f"{name}: shape {tuple(got.shape)} != {tuple(want.shape)}"
    if exact or not got.dtype.is_floating_point:
        bad = (got != want.to(got.dtype))
        assert not bad.any(), f"{name}: {int(bad.sum())} of {bad.numel()} elements differ (bit-exact required)"
        return
    a, b = got.double(), want.double()
    tol = RTOL * torch.maximum(a.abs(), b.abs()) + ATOL
    bad = (a - b).abs() > tol
    if bad.any():
        idx = bad.nonzero()[0].tolist()
        raise AssertionError(f"{name}: {int(bad.sum())} of {bad.numel()} outside 1e-5 rel; first at {idx}: "
                             f"got {got[tuple(idx)].item()!r} want {want[tuple(idx)].item()!r}")


FLOAT_BUFFERS = ("obs_buf", "rew_buf", "commands", "root_states", "dof_state", "last_actions", "last_dof_vel",
                 "last_root_vel", "feet_air_time", "base_lin_vel", "base_ang_vel", "projected_gravity", "env_origins")
EXACT_BUFFERS = ("reset_buf", "time_out_buf", "episode_length_buf")


def compare_env_to_oracle(env, s, tag=""):
    """`env`: this repo's LeggedRobot after a step; `s`: OracleState after the same step."""
    for name in EXACT_BUFFERS:
        assert_close(f"{tag}{name}", getattr(env, name), getattr(s, name), exact=True)
    if hasattr(env, "terrain_levels"):
        assert_close(f"{tag}terrain_levels", env.terrain_levels, s.terrain_levels, exact=True)
    n = int(env.step_counts[0])
    want_ids = s.reset_env_ids.cpu().to(torch.int32)
    assert n == len(want_ids), f"{tag}n_reset {n} != {len(want_ids)}"
    assert_close(f"{tag}reset_ids", env.reset_ids[:n], want_ids, exact=True)
    m = int(env.step_counts[1])
    want_rs = s.resample_env_ids.cpu().to(torch.int32)
    assert m == len(want_rs), f"{tag}n_resample {m} != {len(want_rs)}"
    assert_close(f"{tag}resample_ids", env.resample_ids[:m], want_rs, exact=True)
    if torch.is_tensor(env.measured_heights) and torch.is_tensor(s.measured_heights):
        # int16 * 0.005f: exact iff the height-field indices are exact
        assert_close(f"{tag}measured_heights", env.measured_heights, s.measured_heights, exact=True)
    for name in FLOAT_BUFFERS:
        assert_close(f"{tag}{name}", getattr(env, name).reshape(-1), getattr(s, name).reshape(-1))
    for k in env.episode_sums:
        assert_close(f"{tag}episode_sums[{k}]", env.episode_sums[k], s.episode_sums[k])
    if getattr(env.cfg.rewards, "filtered_feet_contact", False):
        assert_close(f"{tag}last_contacts", env.last_contacts, s.last_contacts, exact=True)
    if n > 0:
        for k, v in s.extras.get("episode", {}).items():
            if k in env.extras["episode"] and torch.is_tensor(env.extras["episode"][k]):
                assert_close(f"{tag}extras[{k}]", env.extras["episode"][k].reshape(1), torch.as_tensor(v).reshape(1).float())


def run_parity(scenario, num_envs, steps, oracle_device="cpu", index_math=None, seed=0, materialize=True, extra=None,
               task_class=None, oracle_extra_terms=None):
    """Steps this repo's env (CUDA kernels, taps) and the oracle on the same frames; compares every step."""
    from legged_gym_b200 import _lgk
    from oracle import legged_oracle as lo
    env = make_env(scenario, num_envs, device="cuda:0", seed=seed, extra=extra, task_class=task_class)
    randomise_env_state(env, seed)
    frames = synthetic.make_frames(env, 4, seed=1234 + seed)
    synthetic.load_frame(env, frames[0])
    p = lo.OracleParams.from_env(env, device=oracle_device)
    p.extra_reward_terms = oracle_extra_terms
    s = lo.OracleState.from_env(env, p, device=oracle_device)
    if index_math is None:
        index_math = _lgk.INDEX_MATH_TORCH_CPU if oracle_device == "cpu" else _lgk.INDEX_MATH_TORCH_CUDA
    total_resets = 0
    for it in range(steps):
        fr = frames[it % len(frames)]
        taps = synthetic.make_taps(env, seed=100 * seed + it)
        synthetic.load_frame(env, fr)
        env.use_rng_taps(taps, index_math=index_math)
        env.step(fr["actions"].clone())
        torch.cuda.synchronize()
        od = {k: v.to(oracle_device) for k, v in fr.items()}
        s.root_states.copy_(od["root_states"]); s.dof_state.copy_(od["dof_state"])
        s.contact_forces.copy_(od["contact_forces"].view_as(s.contact_forces)); s.torques.copy_(od["torques"])
        target = lo.step(p, s, od["actions"].clone(), {k: v.to(oracle_device) for k, v in taps.items()})
        compare_env_to_oracle(env, s, tag=f"[{scenario} step {it}] ")
        assert_close("target_poses", env.target_poses, target)
        total_resets += int(env.step_counts[0])
    return total_resets


# ---------------------------------------------------------------------------------------------
# golden fixtures: outputs of the reference's own code (tests/golden/make_golden.py)
import os
import zlib

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLDEN_OUT = ("obs_buf", "rew_buf", "commands", "root_states", "dof_state", "last_actions", "last_dof_vel",
              "last_root_vel", "feet_air_time", "base_lin_vel", "base_ang_vel", "projected_gravity", "env_origins",
              "target_poses")
GOLDEN_EXACT = ("reset_buf", "time_out_buf", "episode_length_buf")


def load_golden(name):
    d = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    return {k: d[k] for k in d.files}


def _crc(t):
    return zlib.crc32(np.ascontiguousarray(t.detach().cpu().numpy()).tobytes()) & 0xFFFFFFFF


def golden_env(name, g, device):
    """This repo's env built for the fixture's scenario, put into the fixture's initial state."""
    N = int(g["meta_envs"])
    env = make_env(name, N, device=device, seed=int(g["meta_seed"]))
    if int(g["meta_terrain_crc"]) != 0:
        assert _crc(env.height_samples) == int(g["meta_terrain_crc"]), "terrain height field differs from the reference's"
        assert _crc(env.terrain_origins) == int(g["meta_origins_crc"]), "terrain origins differ from the reference's"
    assert list(env.reward_names) == [str(x) for x in g["meta_reward_names"]], "reward evaluation order differs"
    keys = [str(x) for x in g["meta_sum_keys"]]
    assert list(env.episode_sums.keys()) == keys, "episode_sums key order differs"
    T = lambda a: torch.from_numpy(np.asarray(a)).to(device)
    env.episode_length_buf = T(g["init_episode_length_buf"])
    for k in ("commands", "feet_air_time", "last_actions", "last_dof_vel", "env_origins"):
        getattr(env, k).copy_(T(g["init_" + k]))
    for i, k in enumerate(keys):
        env.episode_sums[k].copy_(T(g["init_episode_sums"][i]))
    if g["init_terrain_levels"].size:
        env.terrain_levels.copy_(T(g["init_terrain_levels"]))
        assert torch.equal(env.terrain_types.cpu(), torch.from_numpy(g["init_terrain_types"])), "terrain_types differ"
    env.common_step_counter = int(g["init_common_step_counter"])
    return env, keys


def golden_frame(g, it, device):
    T = lambda a: torch.from_numpy(np.asarray(a)).to(device)
    fr = {k: T(g["in_" + k][it]) for k in ("root_states", "dof_state", "contact_forces", "torques", "actions")}
    taps = {k[4:]: T(g[k][it]) for k in g if k.startswith("tap_")}
    return fr, taps


def compare_to_golden(get, g, it, keys, tag, float_names=GOLDEN_OUT):
    """`get(name)` returns the candidate's buffer; compares against the reference's step-`it` outputs."""
    T = lambda a: torch.from_numpy(np.asarray(a))
    for name in GOLDEN_EXACT:
        assert_close(f"{tag}{name}", get(name).reshape(-1), T(g["out_" + name][it]).reshape(-1), exact=True)
    for name in float_names:
        assert_close(f"{tag}{name}", get(name).reshape(-1), T(g["out_" + name][it]).reshape(-1))
    sums = get("episode_sums")
    for i, k in enumerate(keys):
        assert_close(f"{tag}episode_sums[{k}]", sums[k], T(g["out_episode_sums"][it][i]))
    if g["out_terrain_levels"][it].size:
        assert_close(f"{tag}terrain_levels", get("terrain_levels"), T(g["out_terrain_levels"][it]), exact=True)
    if g["out_measured_heights"][it].size:
        mh = get("measured_heights")
        if torch.is_tensor(mh):
            assert_close(f"{tag}measured_heights", mh, T(g["out_measured_heights"][it]), exact=True)
    want = g["out_reset_ids"][it]
    want = want[want >= 0]
    got = get("reset_ids")
    assert_close(f"{tag}reset_ids", got.to(torch.int64).reshape(-1), T(want), exact=True)
    want = g["out_resample_ids"][it]
    want = want[want >= 0]
    assert_close(f"{tag}resample_ids", get("resample_ids").to(torch.int64).reshape(-1), T(want), exact=True)
    if "out_command_ranges" in g:   # the command curriculum's widened range (legged_robot.py:465-474), exact
        cr = get("command_ranges")
        got_cr = np.array([cr[k] for k in ("lin_vel_x", "lin_vel_y", "ang_vel_yaw", "heading")], np.float64)
        assert np.array_equal(got_cr, g["out_command_ranges"][it]), f"{tag}command_ranges {got_cr} != {g['out_command_ranges'][it]}"
    ex = get("extras")
    if ex is not None:
        row = g["out_extras"][it]
        for i, k in enumerate(keys + ["terrain_level"]):
            if not np.isnan(row[i]):
                name = ("rew_" + k) if i < len(keys) else k
                assert_close(f"{tag}extras[{name}]", torch.as_tensor(ex[name]).float().reshape(1), T(row[i:i + 1]))


def sea_setup(g, device):
    """ANYmal fixture: the LSTM weights the fixture was made with and the torque tolerance.

    torque = out_scale * (b + sum_k w_k h_k) with |h_k| < 1 cancels: its rounding error scales with the
    summands, not with the result, so the absolute floor is 1e-6 of out_scale * (|b| + sum |w_k|), the
    magnitude of the sum's operands (ten times tighter than the north-star relative tolerance on them)."""
    from legged_gym_b200.envs.anymal_c.anymal import random_lstm_state_dict
    sd = random_lstm_state_dict(0)
    atol = 0.1 * RTOL * float(sd["out_scale"].abs().max() * (sd["linear.bias"].abs().sum() + sd["linear.weight"].abs().sum()))
    return sd, atol


def sea_atol_of(sd):
    """torque floor of an LSTM actuator net: 1e-6 of out_scale * (|b| + sum |w_k|) (see sea_setup)"""
    return 0.1 * RTOL * float(sd["out_scale"].abs().max() * (sd["linear.bias"].abs().sum() + sd["linear.weight"].abs().sum()))


def assert_close_abs(name, got, want, atol):
    """1e-5 relative plus an absolute floor `atol` (scalar, or a tensor broadcast over the last dimension)."""
    a, b = got.detach().cpu().double(), want.detach().cpu().double()
    assert a.shape == b.shape, f"{name}: shape {tuple(a.shape)} != {tuple(b.shape)}"
    floor = torch.as_tensor(atol, dtype=torch.double)
    bad = ((a - b).abs() > RTOL * torch.maximum(a.abs(), b.abs()) + floor).reshape(-1)
    if bad.any():
        i = int(bad.nonzero()[0])
        raise AssertionError(f"{name}: {int(bad.sum())} of {bad.numel()} outside 1e-5 rel + {float(floor.max()):.2e}; "
                             f"first at {i}: got {a.reshape(-1)[i].item()!r} want {b.reshape(-1)[i].item()!r}")


def actuator_dvel_atol(state_dict):
    """Absolute floor for the Go1 actuator net's dVel [.., 12].  dVel_j = vel_std_j * (b_j + sum_k w_jk a_k) with
    |a_k| < 1 is a 128-term sum that cancels to ~1e-2 of its operands' magnitude, so two exact-order-agnostic
    fp32 evaluations differ by a few roundings OF THE OPERANDS (torch fp32 is itself up to 1.9e-7 from the fp64
    value of the net output at these weights).  Floor: 4 fp32 roundings (2.4e-7) of |b_j| + sum_k |w_jk|, scaled
    like the output."""
    from legged_gym_b200.envs.go1 import go1
    w3, b3 = state_dict["architecture.6.weight"].double(), state_dict["architecture.6.bias"].double()
    mag = b3.abs() + w3.abs().sum(1)
    return (2.4e-7 * mag * torch.tensor(go1.VEL_STD, dtype=torch.double)).repeat(go1.LEG_NUM)


def replay_golden_oracle(name, device="cpu"):
    """Oracle restatement vs the reference's recorded outputs."""
    from oracle import legged_oracle as lo
    g = load_golden(name)
    env, keys = golden_env(name, g, "cpu")
    fr0, _ = golden_frame(g, 0, "cpu")
    synthetic.load_frame(env, fr0)
    p = lo.OracleParams.from_env(env, device=device)
    s = lo.OracleState.from_env(env, p, device=device)
    act = None
    if "out_dvel" in g:
        from legged_gym_b200.envs.go1.go1 import random_mlp_state_dict
        act = lo.ActuatorOracle(env.num_envs, random_mlp_state_dict(0), device=device)
    sea = None
    if "out_sea_torques" in g:
        sd, sea_atol = sea_setup(g, device)
        sea = lo.SeaLstmOracle(env.num_envs, env.num_actions, sd, device=device)
    n = 0
    for it in range(int(g["meta_steps"])):
        fr, taps = golden_frame(g, it, device)
        s.root_states.copy_(fr["root_states"]); s.dof_state.copy_(fr["dof_state"])
        s.contact_forces.copy_(fr["contact_forces"].view_as(s.contact_forces)); s.torques.copy_(fr["torques"])
        if act is not None:
            a = torch.clip(fr["actions"], -p.clip_actions, p.clip_actions)
            for k in range(p.decimation):
                dv = act.advance(a, s.dof_pos, s.dof_vel)
                assert_close(f"[{name} step {it}.{k}] dVel", dv, torch.from_numpy(g["out_dvel"][it, k]))
        target = lo.step(p, s, fr["actions"].clone(), taps)

        def get(nm):
            if nm == "target_poses":
                return target
            if nm == "reset_ids":
                return s.reset_env_ids
            if nm == "resample_ids":
                return s.resample_env_ids
            if nm == "extras":
                return s.extras.get("episode") if len(s.reset_env_ids) else None
            if nm == "command_ranges":
                return p.command_ranges
            return getattr(s, nm)
        compare_to_golden(get, g, it, keys, f"[oracle {name} step {it}] ")
        n += len(s.reset_env_ids)
        if sea is not None:
            sea.reset(s.reset_env_ids)                       # Anymal.reset_idx, anymal.py:56-60
            a = torch.clip(fr["actions"], -p.clip_actions, p.clip_actions)
            for k in range(p.decimation):
                ds = torch.from_numpy(g["in_sea_dof_state"][it, k]).to(device)
                tq = sea.compute_torques(a, p.action_scale, p.default_dof_pos, ds[..., 0], ds[..., 1])
                assert_close_abs(f"[oracle {name} step {it}.{k}] sea torques", tq,
                                 torch.from_numpy(g["out_sea_torques"][it, k]), sea_atol)
            assert_close(f"[oracle {name} step {it}] sea_hidden_state", sea.sea_hidden_state, torch.from_numpy(g["out_sea_hidden"][it]))
            assert_close(f"[oracle {name} step {it}] sea_cell_state", sea.sea_cell_state, torch.from_numpy(g["out_sea_cell"][it]))
            assert_close(f"[oracle {name} step {it}] sea_input", sea.sea_input, torch.from_numpy(g["out_sea_input"][it]))
    return n


def replay_golden_env(name, index_math=None):
    """This repo's CUDA path (through the C ABI) vs the reference's recorded outputs."""
    from legged_gym_b200 import _lgk
    g = load_golden(name)
    env, keys = golden_env(name, g, "cuda:0")
    n = 0
    for it in range(int(g["meta_steps"])):
        fr, taps = golden_frame(g, it, "cuda:0")
        synthetic.load_frame(env, fr)
        # the fixtures were produced by torch-CPU: its (p / horizontal_scale) is an IEEE division
        env.use_rng_taps(taps, index_math=_lgk.INDEX_MATH_TORCH_CPU if index_math is None else index_math)
        env.step(fr["actions"].clone())
        torch.cuda.synchronize()
        if "out_dvel" in g:
            from legged_gym_b200.envs.go1.go1 import random_mlp_state_dict
            assert_close_abs(f"[{name} step {it}] dVel", env.dVel, torch.from_numpy(g["out_dvel"][it, -1]),
                             actuator_dvel_atol(random_mlp_state_dict(0)))
        nr, ns = int(env.step_counts[0]), int(env.step_counts[1])

        def get(nm):
            if nm == "reset_ids":
                return env.reset_ids[:nr]
            if nm == "resample_ids":
                return env.resample_ids[:ns]
            if nm == "extras":
                return env.extras.get("episode") if nr else None
            return getattr(env, nm)
        compare_to_golden(get, g, it, keys, f"[cuda {name} step {it}] ")
        n += nr
        if "out_sea_torques" in g:
            # torque mode on top of the step: Anymal._compute_torques once per simulator substep; the resets of
            # this step have already cleared the LSTM state of the reset envs (device-side id list)
            _, sea_atol = sea_setup(g, "cpu")
            for k in range(env.cfg.control.decimation):
                ds = torch.from_numpy(g["in_sea_dof_state"][it, k]).to("cuda:0")
                env.dof_state.copy_(ds.view_as(env.dof_state))
                tq = env._compute_torques(env.actions)
                torch.cuda.synchronize()
                assert_close_abs(f"[cuda {name} step {it}.{k}] sea torques", tq,
                                 torch.from_numpy(g["out_sea_torques"][it, k]), sea_atol)
            assert_close(f"[cuda {name} step {it}] sea_hidden_state", env.sea_hidden_state, torch.from_numpy(g["out_sea_hidden"][it]))
            assert_close(f"[cuda {name} step {it}] sea_cell_state", env.sea_cell_state, torch.from_numpy(g["out_sea_cell"][it]))
            assert_close(f"[cuda {name} step {it}] sea_input", env.sea_input, torch.from_numpy(g["out_sea_input"][it]))
    return n
