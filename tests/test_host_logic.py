"""CPU tier: the host-side mirror of the reference interface -- registry, cfg -> constants, reward
registration order and error behaviour, sharding of terrain types, no-CPU-path guarantees."""
import copy
import ctypes as C

import pytest
import torch

import helpers
from legged_gym_b200 import _lgk
from legged_gym_b200.envs import LeggedRobot
from legged_gym_b200.utils.task_registry import task_registry


def test_registry_holds_the_reference_tasks():
    names = {"anymal_c_rough", "anymal_c_flat", "anymal_b", "a1", "cassie", "a1_src", "go1", "aliengo"}
    assert names <= set(task_registry.task_classes)
    assert issubclass(task_registry.get_task_class("go1"), LeggedRobot)
    with pytest.raises(ValueError, match="was not registered"):
        task_registry.make_env("nope", args=object())


def test_vecenv_surface_and_dtypes():
    env = helpers.make_env("go1_rough_full", 48)
    assert (env.num_envs, env.num_obs, env.num_actions, env.num_privileged_obs) == (48, 235, 12, None)
    assert env.obs_buf.shape == (48, 235) and env.obs_buf.dtype == torch.float32
    assert env.rew_buf.shape == (48,) and env.reset_buf.dtype == torch.bool
    assert env.episode_length_buf.dtype == torch.int64 and env.time_out_buf.dtype == torch.bool
    assert env.get_observations() is env.obs_buf and env.get_privileged_observations() is None
    assert float(env.max_episode_length) == 1000.0 and abs(env.dt - 0.02) < 1e-12
    assert env.num_bodies == 17 and env.feet_indices.tolist() == [4, 8, 12, 16]
    assert len(env.penalised_contact_indices) == 8 and len(env.termination_contact_indices) == 1
    assert env.height_samples.dtype == torch.int16 and tuple(env.height_samples.shape) == (1300, 2100)
    assert env.height_points.shape == (48, 187, 3)
    assert env.terrain_levels.dtype == torch.int64 and env.terrain_types.dtype == torch.int64
    # rsl_rl re-binds episode_length_buf (scripts/train.py:43): assignment must write through
    addr = env.episode_length_buf.data_ptr()
    env.episode_length_buf = torch.arange(48)
    assert env.episode_length_buf.data_ptr() == addr and env.episode_length_buf[5] == 5


def test_reward_registration_is_alphabetical_and_scaled_by_dt():
    env = helpers.make_env("go1_rough_full", 16)
    assert env.reward_names == sorted(env.reward_names) and len(env.reward_names) == 20
    assert "termination" in env.reward_scales and "termination" not in env.reward_names
    assert abs(env.reward_scales["tracking_lin_vel"] - 1.0 * env.dt) < 1e-9
    assert list(env.episode_sums.keys()) == list(env.reward_scales.keys())
    flat = helpers.make_env("go1_flat", 16)
    assert flat.reward_names == ["action_rate", "ang_vel_xy", "collision", "dof_acc", "dof_pos_limits",
                                 "feet_air_time", "lin_vel_z", "torques", "tracking_ang_vel", "tracking_lin_vel"]
    cassie = helpers.make_env("cassie_flat", 16)
    assert "no_fly" in cassie.reward_names and cassie.num_obs == 169 and cassie.num_bodies == 13


def test_unknown_reward_and_control_type_raise_like_the_reference():
    def stumble(cfg):
        cfg.rewards.scales.feet_stumble = -1.0       # no _reward_feet_stumble exists (legged_robot.py:593)
    with pytest.raises(AttributeError, match="_reward_feet_stumble"):
        helpers.make_env("go1_flat", 8, extra=stumble)
    env = helpers.make_env("go1_flat", 8)
    env.cfg.control.control_type = "X"
    with pytest.raises(NameError, match="Unknown controller type"):
        env.build_consts()


def test_noise_scale_vector():
    env = helpers.make_env("go1_rough_full", 8)
    v = env.noise_scale_vec
    assert torch.allclose(v[:3], torch.full((3,), 0.1 * 2.0)) and torch.allclose(v[3:6], torch.full((3,), 0.2 * 0.25))
    assert torch.allclose(v[6:9], torch.full((3,), 0.05)) and (v[9:12] == 0).all() and (v[36:48] == 0).all()
    assert torch.allclose(v[12:24], torch.full((12,), 0.01)) and torch.allclose(v[24:36], torch.full((12,), 1.5 * 0.05))
    assert torch.allclose(v[48:], torch.full((187,), 0.1 * 5.0))


def test_constant_block_mirrors_cfg():
    env = helpers.make_env("go1_rough_full", 64)
    c = env.build_consts()
    assert c.struct_size == C.sizeof(_lgk.LgkConsts)
    assert (c.num_envs, c.num_dof, c.num_bodies, c.num_feet, c.num_pen, c.num_term) == (64, 12, 17, 4, 8, 1)
    assert (c.num_height_points, c.num_obs, c.num_rewards) == (187, 235, 20)
    assert c.flags & _lgk.F_HAS_TERMINATION and c.flags & _lgk.F_ONLY_POSITIVE and c.flags & _lgk.F_TERRAIN_CURRICULUM
    assert c.flags & _lgk.F_HEADING_COMMAND and c.flags & _lgk.F_CUSTOM_ORIGINS and not c.flags & _lgk.F_TERRAIN_PLANE
    assert (c.max_episode_length, c.resample_interval, c.push_interval) == (1000, 500, 750)
    assert [c.rewards[i].term for i in range(20)] == [_lgk.REWARD_ID[n] for n in env.reward_names]
    assert abs(c.rewards[0].scale - env.reward_scales["action_rate"]) < 1e-10      # fp32 of the python float
    assert (c.terrain_rows, c.terrain_cols, c.terrain_pitch) == (1300, 2100, 2104)
    assert list(c.hip_idx) == [0, 3, 6, 9]
    cas = helpers.make_env("cassie_flat", 8).build_consts()
    assert cas.flags & _lgk.F_TERRAIN_PLANE and cas.num_height_points == 121 and cas.num_feet == 2


def test_sharded_terrain_types_follow_the_global_env_index():
    def shard(rank):
        def f(cfg):
            cfg.env.shard_rank, cfg.env.shard_world = rank, 4
        return f
    whole = helpers.make_env("go1_rough_full", 256)
    parts = [helpers.make_env("go1_rough_full", 64, extra=shard(r)) for r in range(4)]
    assert torch.equal(torch.cat([p.terrain_types for p in parts]), whole.terrain_types)
    assert [p.env_id_offset for p in parts] == [0, 64, 128, 192]
    seeds = {p.build_consts().seed for p in parts}
    assert len(seeds) == 4         # independent Philox streams per rank


def test_there_is_no_cpu_step_path():
    env = helpers.make_env("go1_flat", 8, device="cpu")
    with pytest.raises(RuntimeError, match="no CPU path"):
        env.step(torch.zeros(8, 12))
    with pytest.raises(RuntimeError, match="no CPU path"):
        env.reset()


def test_fused_hooks_cannot_be_overridden_silently():
    class Bad(LeggedRobot):
        def _resample_commands(self, env_ids):      # runs inside the fused kernel: an override would be ignored
            pass
    cfg = copy.deepcopy(task_registry.env_cfgs["a1"])
    cfg.env.num_envs = 4
    from legged_gym_b200.sim import gymapi, gymutil
    from legged_gym_b200.utils.helpers import class_to_dict
    sp = gymapi.SimParams()
    gymutil.parse_sim_config(class_to_dict(cfg.sim), sp)
    with pytest.raises(NotImplementedError, match="_resample_commands"):
        Bad(cfg=cfg, sim_params=sp, physics_engine=gymapi.SIM_PHYSX, sim_device="cpu", headless=True)


def test_product_code_never_imports_the_oracle():
    import os
    import re
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "legged_gym_b200")
    for dp, _, files in os.walk(root):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f"{f} imports oracle/"


@pytest.mark.parametrize("scenario", ["go1_rough_full", "go1_rough_actuator", "anymal_c_flat"])
def test_env_state_checkpoint_roundtrip(scenario, tmp_path):
    """state_dict() -> torch.save -> load_state_dict() restores every state buffer in place."""
    env = helpers.make_env(scenario, 24, seed=2)
    helpers.randomise_env_state(env, 2)
    env.command_ranges["lin_vel_x"][1] = 1.5          # as the command curriculum would have left it
    env.common_step_counter = 4711
    extra = env._extra_state_tensors()
    for t in extra.values():
        t.copy_(torch.randn(t.shape))
    sd = env.state_dict()
    path = tmp_path / "env_state.pt"
    torch.save(sd, path)
    want = {n: t.clone() for n, t in sd["tensors"].items()}
    other = helpers.make_env(scenario, 24, seed=2)
    addr = other.commands.data_ptr()
    other.load_state_dict(torch.load(path, weights_only=False))
    assert other.common_step_counter == 4711 and other.command_ranges["lin_vel_x"][1] == 1.5
    assert other.commands.data_ptr() == addr          # restored in place: kernel pointers stay valid
    got = other.state_dict()["tensors"]
    assert set(got) == set(want) and len(want) >= 18 + len(extra)
    for n in want:
        assert torch.equal(got[n], want[n]), n
    with pytest.raises(ValueError, match="num_envs"):
        helpers.make_env(scenario, 8).load_state_dict(sd)


def test_terrain_file_roundtrip(tmp_path):
    """SURVEY 8(f) rank 3 (format half): a generated terrain saved to disk and loaded through
    cfg.terrain.terrain_file gives the same height field, origins and env placement without touching numpy's RNG."""
    a = helpers.make_env("go1_rough_full", 16, seed=3)
    path = str(tmp_path / "terrain.npz")
    a.terrain.save(path)

    def from_file(cfg):
        cfg.terrain.terrain_file = path
    b = helpers.make_env("go1_rough_full", 16, seed=4, extra=from_file)   # another numpy stream: the field comes from the file
    assert torch.equal(a.height_samples, b.height_samples) and torch.equal(a.terrain_origins, b.terrain_origins)

    def wrong_scale(cfg):
        cfg.terrain.terrain_file = path
        cfg.terrain.vertical_scale = 0.01
    with pytest.raises(ValueError, match="vertical_scale"):
        helpers.make_env("go1_rough_full", 16, extra=wrong_scale)


def test_privileged_observations_buffer_like_the_reference():
    """base_task.py:75-78: with cfg.env.num_privileged_obs set the buffer exists (zeros; no shipped task writes
    it) and get_privileged_observations() returns it; None otherwise."""
    def priv(cfg):
        cfg.env.num_privileged_obs = 17
    env = helpers.make_env("go1_flat", 6, extra=priv)
    assert env.num_privileged_obs == 17 and env.get_privileged_observations() is env.privileged_obs_buf
    assert env.privileged_obs_buf.shape == (6, 17) and float(env.privileged_obs_buf.abs().sum()) == 0.0
    assert helpers.make_env("go1_flat", 6).get_privileged_observations() is None


def test_reset_idx_override_without_the_fused_hook_fails_at_construction():
    """step() resets inside the fused kernel and never calls reset_idx(): an override that clears extra per-env
    state must come with _after_fused_resets(), or construction fails."""
    import helpers
    from legged_gym_b200.envs.base.legged_robot import LeggedRobot

    class Bad(LeggedRobot):
        def reset_idx(self, env_ids):
            super().reset_idx(env_ids)

    class Good(Bad):
        def _after_fused_resets(self):
            pass
    with pytest.raises(NotImplementedError, match="_after_fused_resets"):
        helpers.make_env("go1_flat", 8, device="cpu", task_class=Bad)
    helpers.make_env("go1_flat", 8, device="cpu", task_class=Good)


def test_legged_gym_import_path_alias():
    """User code written against the reference's module paths (envs/__init__.py:33-59) runs unchanged with
    <repo>/compat on the path: legged_gym.<x> IS legged_gym_b200.<x>.  In a subprocess: this process may hold the
    reference itself under that name (oracle)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = ("from legged_gym.envs.base.legged_robot import LeggedRobot\n"
            "from legged_gym.envs.base.legged_robot_config import LeggedRobotCfg\n"
            "from legged_gym.utils.task_registry import task_registry\n"
            "from legged_gym.utils import helpers\n"
            "from legged_gym import LEGGED_GYM_ROOT_DIR\n"
            "import legged_gym.envs, legged_gym_b200.envs.base.legged_robot as real\n"
            "assert LeggedRobot is real.LeggedRobot\n"
            "assert 'go1' in task_registry.task_classes and 'anymal_c_rough' in task_registry.task_classes\n"
            "class MyRobot(LeggedRobot):\n    def _reward_custom(self):\n        return self.base_lin_vel[:, 0]\n"
            "print('ok')\n")
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([os.path.join(root, "compat"), root]))
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and out.stdout.strip().endswith("ok"), out.stderr[-2000:]


def _emulate_terrain_recipes(ter, subs, draws, lin):
    """numpy restatement of lgk_terrain.cu's per-cell formulas (so that the recipes and the formulas can be checked
    against the host generators without a GPU)."""
    import numpy as np
    from legged_gym_b200 import _lgk
    W = L = ter.width_per_env_pixels
    field = np.zeros((ter.tot_rows, ter.tot_cols), np.int16)
    X, Y = np.meshgrid(np.arange(W), np.arange(L), indexing="ij")
    for t in subs:
        p, d = list(t.p), np.array(draws[t.draws_off:] + [0], np.int64)
        if t.type in (_lgk.TER_SLOPE, _lgk.TER_SLOPE_NOISE):
            cx, cy = W // 2, L // 2
            h = (p[0] * ((cx - np.abs(cx - X)) / cx) * ((cy - np.abs(cy - Y)) / cy)).astype(np.int16).astype(np.int64)
            h = np.clip(h, p[1], p[2])
            if t.type == _lgk.TER_SLOPE_NOISE:
                nx, ny = p[3], p[4]
                gx, gy = lin[:W][X], lin[W:][Y]
                x0 = np.clip(np.floor(gx).astype(int), 0, nx - 2); y0 = np.clip(np.floor(gy).astype(int), 0, ny - 2)
                fx, fy = gx - x0, gy - y0
                x1, y1 = np.minimum(x0 + 1, nx - 1), np.minimum(y0 + 1, ny - 1)
                c = d[:nx * ny].reshape(nx, ny).astype(np.float64)
                f = c[x0, y0] * (1 - fx) * (1 - fy) + c[x1, y0] * fx * (1 - fy) + c[x0, y1] * (1 - fx) * fy + c[x1, y1] * fx * fy
                h = h + np.rint(f).astype(np.int16)
        elif t.type == _lgk.TER_STAIRS:
            sw = p[0]
            r = np.minimum(np.minimum(X // sw, (W - 1 - X) // sw), np.minimum(Y // sw, (L - 1 - Y) // sw))
            h = p[1] * np.minimum(r, p[2])
        elif t.type == _lgk.TER_OBSTACLES:
            h = np.zeros((W, L), np.int64)
            for k in range(p[0]):
                i, j, w, l, hh = d[5 * k:5 * k + 5]
                h[(X >= i) & (X < i + w) & (Y >= j) & (Y < j + l)] = hh
            h[(X >= p[1]) & (X < p[2]) & (Y >= p[3]) & (Y < p[4])] = 0
        elif t.type == _lgk.TER_STONES:
            ss, pitch = p[0], p[0] + p[1]
            sa, sb = X // pitch, Y // pitch
            on = (X - sa * pitch < ss) & (Y - sb * pitch < ss)
            h = np.where(on, d[np.minimum(sa * p[3] + sb, len(d) - 1)], p[2])
            h[(X >= p[4]) & (X < p[5]) & (Y >= p[6]) & (Y < p[7])] = 0
        elif t.type == _lgk.TER_GAP:
            cx, cy = L // 2, W // 2
            h = np.zeros((W, L), np.int64)
            h[(X >= cx - p[0]) & (X < cx + p[0]) & (Y >= cy - p[1]) & (Y < cy + p[1])] = -1000
            h[(X >= cx - p[2]) & (X < cx + p[2]) & (Y >= cy - p[3]) & (Y < cy + p[3])] = 0
        elif t.type == _lgk.TER_PIT:
            cx, cy = L // 2, W // 2
            h = np.zeros((W, L), np.int64)
            h[(X >= cx - p[1]) & (X < cx + p[1]) & (Y >= cy - p[1]) & (Y < cy + p[1])] = -p[0]
        else:
            h = np.zeros((W, L), np.int64)
        x0 = ter.border + t.row * ter.length_per_env_pixels
        y0 = ter.border + t.col * W
        field[x0:x0 + W, y0:y0 + L] = h.astype(np.int16)
    return field


@pytest.mark.parametrize("curriculum", [True, False])
def test_device_terrain_recipes_reproduce_the_host_generator(curriculum):
    """The recipes lgk_generate_terrain consumes (Terrain._recipe: parameters + resolved np.random draws in the
    generators' own order) describe exactly the field the host generators build, for every sub-terrain type
    (slope, slope + noise, stairs up / down, obstacles, stepping stones, gap, pit) -- checked through a numpy
    restatement of the kernel's per-cell formulas; the np.random stream ends in the same state."""
    import copy
    import numpy as np
    from legged_gym_b200.envs.base.legged_robot_config import LeggedRobotCfg
    from legged_gym_b200.utils.terrain import Terrain
    cfg = type("TerrainCfg", (LeggedRobotCfg.terrain,), {})()      # an instance: the class attributes stay untouched
    cfg.mesh_type, cfg.curriculum, cfg.selected = "heightfield", curriculum, False
    cfg.num_rows, cfg.num_cols = 5, 14
    cfg.terrain_proportions = [0.1, 0.1, 0.2, 0.2, 0.1, 0.1, 0.1]   # + pit: every branch of make_terrain
    np.random.seed(7)
    host = Terrain(copy.copy(cfg), 16)
    end_state_host = np.random.get_state()[1].copy()
    np.random.seed(7)
    ter = copy.copy(host)          # geometry and proportions; the field itself is rebuilt from the recipes below
    subs, draws = [], []
    if curriculum:
        for j in range(cfg.num_cols):
            for i in range(cfg.num_rows):
                subs.append(ter._recipe(j / cfg.num_cols + 0.001, i / cfg.num_rows, i, j, draws))
    else:
        for k in range(cfg.num_rows * cfg.num_cols):
            i, j = np.unravel_index(k, (cfg.num_rows, cfg.num_cols))
            choice = np.random.uniform(0, 1)
            difficulty = np.random.choice([0.5, 0.75, 0.9])
            subs.append(ter._recipe(choice, difficulty, int(i), int(j), draws))
    assert (np.random.get_state()[1] == end_state_host).all(), "the recipes consume the np.random stream differently"
    W = ter.width_per_env_pixels
    nx = int(W * cfg.horizontal_scale / 0.2)
    lin = np.concatenate([np.linspace(0, nx - 1, W), np.linspace(0, nx - 1, W)])
    field = _emulate_terrain_recipes(ter, subs, draws, lin)
    assert len({t.type for t in subs}) >= (7 if curriculum else 4)
    assert field.shape == host.height_field_raw.shape
    bad = np.argwhere(field != host.height_field_raw)
    assert bad.size == 0, f"{len(bad)} cells differ, first at {bad[0]}"

