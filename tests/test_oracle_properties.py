"""CPU tier: edge cases of the path, checked on the oracle (the behaviour the CUDA path is held to):
out-of-range height indices are clipped, zero-norm commands are zeroed, an empty reset set is a no-op,
the terrain level wraps at the top, truncation (not floor) of negative coordinates."""
import torch

import helpers
from legged_gym_b200.utils import synthetic
from oracle import legged_oracle as lo


def _setup(name, n, seed=0):
    env = helpers.make_env(name, n, seed=seed)
    helpers.randomise_env_state(env, seed)
    fr = synthetic.make_frames(env, 1, seed=7)[0]
    synthetic.load_frame(env, fr)
    p = lo.OracleParams.from_env(env)
    s = lo.OracleState.from_env(env, p)
    return env, p, s, fr


def test_height_indices_clip_outside_the_terrain():
    env, p, s, fr = _setup("go1_rough_full", 16)
    s.root_states[:, 0] = torch.linspace(-60., 200., 16)      # far outside [-25, 105] m
    s.root_states[:, 1] = torch.linspace(300., -80., 16)
    h = lo.get_heights(p, s)
    idx = s.height_indices
    assert int(idx[..., 0].min()) == 0 and int(idx[..., 0].max()) == p.height_samples.shape[0] - 2
    assert int(idx[..., 1].min()) == 0 and int(idx[..., 1].max()) == p.height_samples.shape[1] - 2
    assert torch.isfinite(h).all()
    # trunc toward zero: coordinates in (-1, 0) cells land on index 0 without clipping being needed
    s.root_states[:, :2] = -25.05
    s.root_states[:, 3:7] = torch.tensor([0., 0., 0., 1.])
    lo.get_heights(p, s)
    assert (s.height_indices >= 0).all()


def test_plane_terrain_measures_zero_heights():
    env, p, s, fr = _setup("cassie_flat", 8)
    assert lo.get_heights(p, s).abs().sum() == 0 and lo.get_heights(p, s).shape == (8, 121)


def test_small_commands_are_zeroed_and_empty_reset_is_a_noop():
    env, p, s, fr = _setup("go1_flat", 32)
    u = torch.full((32, 3), 0.5)           # middle of [-1, 1] -> exactly 0 -> norm 0 <= 0.2
    lo.resample_commands(p, s, torch.arange(32), u)
    assert (s.commands[:, :2] == 0).all()
    before = {k: getattr(s, k).clone() for k in ("root_states", "dof_state", "commands", "episode_length_buf")}
    s.extras.clear()
    lo.reset_idx(p, s, torch.zeros(0, dtype=torch.long), synthetic.make_taps(env, 1))
    assert all(torch.equal(getattr(s, k), v) for k, v in before.items()) and "episode" not in s.extras


def test_terrain_level_wraps_at_the_top_and_clamps_at_zero():
    env, p, s, fr = _setup("go1_rough_full", 8)
    taps = synthetic.make_taps(env, 3)
    ids = torch.arange(8)
    s.terrain_levels[:] = p.max_terrain_level - 1
    s.root_states[:, :2] = s.env_origins[:, :2] + 5.0          # walked > env_length / 2 -> move up
    lo.update_terrain_curriculum(p, s, ids, taps)
    assert torch.equal(s.terrain_levels, taps["level"])          # wrapped to the random level
    s.terrain_levels[:] = 0
    s.root_states[:, :2] = s.env_origins[:, :2]                 # did not move -> move down, clamped at 0
    s.commands[:, :2] = 1.0
    lo.update_terrain_curriculum(p, s, ids, taps)
    assert (s.terrain_levels == 0).all()
    assert torch.equal(s.env_origins, p.terrain_origins[s.terrain_levels, p.terrain_types])


def test_step_ordering_rules():
    """SURVEY A-2: rewards see pre-reset state; observations mix pre-reset velocities with post-reset joints."""
    env, p, s, fr = _setup("go1_rough_full", 64)
    s.episode_length_buf[:4] = int(p.max_episode_length)        # time-outs
    taps = synthetic.make_taps(env, 9)
    lo.step(p, s, fr["actions"].clone(), taps)
    ids = s.reset_env_ids
    assert set(range(4)) <= set(ids.tolist()) and s.time_out_buf[:4].all()
    assert (s.episode_length_buf[ids] == 0).all() and (s.feet_air_time[ids] == 0).all()
    assert (s.dof_vel[ids] == 0).all() and (s.last_dof_vel[ids] == 0).all()
    assert torch.equal(s.last_actions, s.actions)
    assert (s.obs_buf.abs() <= p.clip_observations).all()
    for k in s.episode_sums:
        assert (s.episode_sums[k][ids] == 0).all()
