"""The five workload configurations BASELINE.json names, as cfg mutators.

A mutator takes a task cfg *instance* (this package's or the reference's -- the attribute names
are the same) and edits it in place.  SURVEY.md section 8(d) defines the configurations.
"""
from __future__ import annotations

# scales added on top of the shipped Go1 set to enable every reward term ("full reward set")
FULL_REWARD_EXTRA = dict(
    orientation=-0.2, dof_vel=-1.e-4, base_height=-0.5, stand_still=-0.01, energy=-5.e-5, hip_motion=-0.02,
    dof_vel_limits=-0.1, torque_limits=-0.01, feet_contact_forces=-0.01, stumble=-0.1, termination=-1.0,
)


def _rough(cfg):
    cfg.terrain.mesh_type = 'trimesh'
    cfg.terrain.measure_heights = True
    cfg.terrain.curriculum = True
    cfg.env.num_observations = 235


def go1_flat(cfg):
    """config 1: Go1 as shipped (plane, blind, 48 obs, 10 reward terms), PD position targets only."""
    cfg.control.use_actuator_network = False


def go1_rough_full(cfg):
    """config 2: rough terrain, 187-point scan, all 20 reward terms + termination."""
    _rough(cfg)
    cfg.control.use_actuator_network = False
    for k, v in FULL_REWARD_EXTRA.items():
        setattr(cfg.rewards.scales, k, v)
    cfg.rewards.only_positive_rewards = True


def go1_rough_default(cfg):
    """rough terrain + scan with the shipped 10-term reward set."""
    _rough(cfg)
    cfg.control.use_actuator_network = False


def go1_rough_actuator(cfg):
    """config 3: config 2 + the actuator-net MLP every simulator substep."""
    go1_rough_full(cfg)
    cfg.control.use_actuator_network = True


def go1_flat_cmd_curriculum(cfg):
    """config 1 with the command curriculum on (legged_robot.py:465-474): the fixture's step counter is set so
    that the gated step (common_step_counter % max_episode_length == 0) falls on its third step."""
    cfg.control.use_actuator_network = False
    cfg.commands.curriculum = True
    cfg.commands.max_curriculum = 2.0   # shipped: 1.0 = the shipped lin_vel_x range, i.e. a curriculum that cannot move


def cassie_flat(cfg):
    """config 4: Cassie as shipped (plane, 121 zero heights, 169 obs, 10 terms + termination)."""


def anymal_c_flat(cfg):
    """ANYmal C as shipped (plane, blind, 48 obs) with the LSTM actuator net."""


def anymal_c_rough(cfg):
    """ANYmal C rough as shipped: trimesh terrain, 187-point scan, terrain curriculum, LSTM actuator net."""


def a1_variant(cfg):
    """A1 with the switches the other scenarios leave at their defaults flipped: heightfield terrain without
    curriculum (randomised sub-terrains, custom origins), yaw-rate instead of heading commands, no
    observation noise, no pushes, negative total rewards kept."""
    cfg.terrain.mesh_type = 'heightfield'
    cfg.terrain.curriculum = False
    cfg.terrain.measure_heights = True
    cfg.env.num_observations = 235
    cfg.commands.heading_command = False
    cfg.noise.add_noise = False
    cfg.domain_rand.push_robots = False
    cfg.rewards.only_positive_rewards = False


def fixture_step_counter(env) -> int:
    """common_step_counter a fixture / parity run starts from: a push (:350) on its third step and, with the command
    curriculum on, the curriculum's gated step (:172) on that same step."""
    import math
    push = int(env.cfg.domain_rand.push_interval)
    if getattr(env.cfg.commands, "curriculum", False):
        return math.lcm(push, int(env.max_episode_length)) - 3
    return push - 3


def fixture_bias_tracking_sums(env, sums) -> None:
    """With the command curriculum on, lift the tracking_lin_vel sums to the curriculum's threshold (:468: mean over
    the reset envs / max_episode_length > 0.8 * scale) so that the decision hinges on the values, not on the sign."""
    if getattr(env.cfg.commands, "curriculum", False) and "tracking_lin_vel" in sums:
        thr = 0.8 * float(env.reward_scales["tracking_lin_vel"]) * float(env.max_episode_length)
        sums["tracking_lin_vel"] += thr + 1.0


SCENARIOS = {
    "go1_flat": ("go1", go1_flat),
    "go1_flat_cmd_curriculum": ("go1", go1_flat_cmd_curriculum),
    "go1_rough_full": ("go1", go1_rough_full),
    "go1_rough_default": ("go1", go1_rough_default),
    "go1_rough_actuator": ("go1", go1_rough_actuator),
    "cassie_flat": ("cassie", cassie_flat),
    "anymal_c_flat": ("anymal_c_flat", anymal_c_flat),
    "anymal_c_rough": ("anymal_c_rough", anymal_c_rough),
    "a1_variant": ("a1", a1_variant),
}


def _as_shipped(cfg):
    """no change: the task exactly as registered"""


# every registered task exactly as shipped (reference envs/__init__.py:52-59); parity-tested against the oracle
# on the GPU, without golden fixtures of their own
SHIPPED_TASKS = {name + "_shipped": (name, _as_shipped) for name in
                 ("anymal_c_rough", "anymal_c_flat", "anymal_b", "a1", "cassie", "a1_src", "go1", "aliengo")}


def algorithmic_bytes_per_env_step(D, B, F, R, O):
    """Compulsory HBM bytes of one post_physics_step per env: reads root 13 + dof_state 2D
    + contacts 3B + torques D + actions D + last_actions D + last_dof_vel D + commands 4 + feet_air_time F
    + episode_sums R floats + episode_length int64; writes obs O + rew 1 + body-frame 9 + last_actions D
    + last_dof_vel D + last_root_vel 6 + feet_air_time F + episode_sums R + commands 4 floats + reset/time_out
    bools + episode_length int64."""
    reads = 4 * (17 + 6 * D + 3 * B + F + R) + 8
    writes = 4 * (O + 20 + 2 * D + F + R) + 10
    return reads + writes
