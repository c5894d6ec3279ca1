"""CPU restatement of the reference's per-env step path -- TEST INFRASTRUCTURE (see oracle/__init__.py).

Plain torch fp32, op for op in the order the reference issues them, so that on a given torch device
the results are bit-identical to the reference's own code (validated in this container against
/root/reference run through ``reference_loader``; fixtures in tests/golden).  It can run on CUDA
torch too (``device='cuda'``): that is the reference's real execution mode and the oracle of record
for the height-field indices.

Differences from the reference, all outside the arithmetic:
  * state lives in an ``OracleState`` object instead of ``self``; constants in ``OracleParams``;
  * random numbers come from explicit per-env tap tensors (row i belongs to env i) instead of the
    global generator -- ``reference_loader.TapHarness`` feeds the reference the same numbers;
  * simulator calls are dropped (they carry no arithmetic).

Every function cites the reference lines it follows; paths are relative to
/root/reference/legged_gym/.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np
import torch


# ----------------------------------------------------------------------------------------------
# isaacgym.torch_utils restatements (closed dependency)
def quat_rotate_inverse(q, v):
    n = q.shape[0]
    q_w = q[:, -1]
    q_vec = q[:, :3]
    a = v * (2.0 * q_w ** 2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * torch.bmm(q_vec.view(n, 1, 3), v.view(n, 3, 1)).squeeze(-1) * 2.0
    return a - b + c


def quat_apply(a, b):
    shape = b.shape
    a = a.reshape(-1, 4)
    b = b.reshape(-1, 3)
    xyz = a[:, :3]
    t = xyz.cross(b, dim=-1) * 2
    return (b + a[:, 3:] * t + xyz.cross(t, dim=-1)).view(shape)


def normalize(x, eps: float = 1e-9):
    return x / x.norm(p=2, dim=-1).clamp(min=eps, max=None).unsqueeze(-1)


def quat_apply_yaw(quat, vec):          # utils/math.py:38-42
    quat_yaw = quat.clone().view(-1, 4)
    quat_yaw[:, :2] = 0.
    quat_yaw = normalize(quat_yaw)
    return quat_apply(quat_yaw, vec)


def wrap_to_pi(angles):                 # utils/math.py:45-48
    angles %= 2 * np.pi
    angles -= 2 * np.pi * (angles > np.pi)
    return angles


def rand_affine(lower, upper, u):       # torch_rand_float with the uniform supplied (SURVEY A-1)
    return (upper - lower) * u + lower


# ----------------------------------------------------------------------------------------------
@dataclass
class OracleParams:
    """Everything _parse_cfg / _init_buffers / _prepare_reward_function derive from the cfg
    (envs/base/legged_robot.py:503-598,769-779), as plain data."""
    num_envs: int
    num_dof: int
    num_bodies: int
    num_obs: int
    dt: float
    sim_dt: float
    decimation: int
    max_episode_length: float
    max_episode_length_s: float
    resampling_interval: int
    push_interval: float
    push_robots: bool
    max_push_vel_xy: float
    heading_command: bool
    command_ranges: Dict[str, List[float]]
    commands_curriculum: bool
    max_curriculum: float
    measure_heights: bool
    mesh_type: str
    terrain_curriculum: bool
    custom_origins: bool
    border_size: float
    horizontal_scale: float
    vertical_scale: float
    env_length: float
    max_terrain_level: int
    add_noise: bool
    clip_actions: float
    clip_observations: float
    obs_scales: Dict[str, float]
    action_scale: float
    control_type: str
    only_positive_rewards: bool
    filtered_feet_contact: bool
    tracking_sigma: float
    soft_dof_vel_limit: float
    soft_torque_limit: float
    base_height_target: float
    max_contact_force: float
    send_timeouts: bool
    reward_scales: Dict[str, float]          # non-zero, already multiplied by dt, reference key order
    reward_names: List[str]                  # evaluation order (termination excluded)
    # tensors
    p_gains: torch.Tensor = None
    d_gains: torch.Tensor = None
    torque_limits: torch.Tensor = None
    default_dof_pos: torch.Tensor = None     # [1, D]
    dof_pos_limits: torch.Tensor = None      # [D, 2]
    dof_vel_limits: torch.Tensor = None
    base_init_state: torch.Tensor = None     # [13]
    noise_scale_vec: torch.Tensor = None
    commands_scale: torch.Tensor = None
    height_points: torch.Tensor = None       # [H, 3]
    height_samples: Optional[torch.Tensor] = None   # int16 [rows, cols]
    terrain_origins: Optional[torch.Tensor] = None
    terrain_types: Optional[torch.Tensor] = None
    feet_indices: torch.Tensor = None
    penalised_contact_indices: torch.Tensor = None
    termination_contact_indices: torch.Tensor = None
    # reward terms a subclass adds as python methods (legged_robot.py:588-593 finds them by getattr): name ->
    # callable(p, s) evaluated at the reference's point in compute_reward()
    extra_reward_terms: Optional[dict] = None

    @staticmethod
    def from_env(env, device=None) -> "OracleParams":
        """Snapshot of an env's constants (works on this repo's LeggedRobot and on the reference's)."""
        cfg = env.cfg
        dev = device or env.device

        def T(x):
            return None if x is None else x.detach().clone().to(dev)
        hp = getattr(env, "height_points", None)
        return OracleParams(
            num_envs=env.num_envs, num_dof=env.num_dof, num_bodies=env.num_bodies, num_obs=env.num_obs,
            dt=env.dt, sim_dt=env.sim_params.dt, decimation=cfg.control.decimation,
            max_episode_length=float(env.max_episode_length), max_episode_length_s=env.max_episode_length_s,
            resampling_interval=int(cfg.commands.resampling_time / env.dt),
            push_interval=float(cfg.domain_rand.push_interval), push_robots=cfg.domain_rand.push_robots,
            max_push_vel_xy=cfg.domain_rand.max_push_vel_xy, heading_command=cfg.commands.heading_command,
            command_ranges={k: list(v) for k, v in env.command_ranges.items()},
            commands_curriculum=cfg.commands.curriculum, max_curriculum=cfg.commands.max_curriculum,
            measure_heights=cfg.terrain.measure_heights, mesh_type=cfg.terrain.mesh_type,
            terrain_curriculum=cfg.terrain.curriculum, custom_origins=env.custom_origins,
            border_size=cfg.terrain.border_size, horizontal_scale=cfg.terrain.horizontal_scale,
            vertical_scale=cfg.terrain.vertical_scale,
            env_length=getattr(getattr(env, "terrain", None), "env_length", 0.0),
            max_terrain_level=getattr(env, "max_terrain_level", 0), add_noise=env.add_noise,
            clip_actions=cfg.normalization.clip_actions, clip_observations=cfg.normalization.clip_observations,
            obs_scales={k: getattr(env.obs_scales, k) for k in
                        ("lin_vel", "ang_vel", "dof_pos", "dof_vel", "height_measurements")},
            action_scale=cfg.control.action_scale, control_type=cfg.control.control_type,
            only_positive_rewards=cfg.rewards.only_positive_rewards, tracking_sigma=cfg.rewards.tracking_sigma,
            filtered_feet_contact=bool(getattr(cfg.rewards, "filtered_feet_contact", False)),
            soft_dof_vel_limit=cfg.rewards.soft_dof_vel_limit, soft_torque_limit=cfg.rewards.soft_torque_limit,
            base_height_target=cfg.rewards.base_height_target, max_contact_force=cfg.rewards.max_contact_force,
            send_timeouts=cfg.env.send_timeouts, reward_scales=dict(env.reward_scales),
            reward_names=list(env.reward_names),
            p_gains=T(env.p_gains), d_gains=T(env.d_gains), torque_limits=T(env.torque_limits),
            default_dof_pos=T(env.default_dof_pos), dof_pos_limits=T(env.dof_pos_limits),
            dof_vel_limits=T(env.dof_vel_limits), base_init_state=T(env.base_init_state),
            noise_scale_vec=T(env.noise_scale_vec), commands_scale=T(env.commands_scale),
            height_points=None if hp is None else hp[0].detach().clone().to(dev),
            height_samples=T(getattr(env, "height_samples", None)),
            terrain_origins=T(getattr(env, "terrain_origins", None)),
            terrain_types=T(getattr(env, "terrain_types", None)),
            feet_indices=T(env.feet_indices), penalised_contact_indices=T(env.penalised_contact_indices),
            termination_contact_indices=T(env.termination_contact_indices),
        )

    def to(self, device) -> "OracleParams":
        import copy
        out = copy.copy(self)
        for k, v in vars(self).items():
            if torch.is_tensor(v):
                setattr(out, k, v.to(device))
        out.command_ranges = {k: list(v) for k, v in self.command_ranges.items()}
        return out


STATE_TENSORS = ("root_states", "dof_state", "contact_forces", "torques", "actions", "last_actions", "last_dof_vel",
                 "last_root_vel", "commands", "feet_air_time", "episode_length_buf", "env_origins",
                 "terrain_levels", "obs_buf", "rew_buf", "reset_buf", "time_out_buf", "base_lin_vel",
                 "base_ang_vel", "projected_gravity")


class OracleState:
    """Mutable per-env state: the reference's buffers (base_task.py:70-78, legged_robot.py:503-572)."""

    def __init__(self, p: OracleParams, device="cpu"):
        N, D, B = p.num_envs, p.num_dof, p.num_bodies
        f = dict(dtype=torch.float, device=device)
        self.root_states = torch.zeros(N, 13, **f)
        self.root_states[:, 6] = 1.0
        self.dof_state = torch.zeros(N * D, 2, **f)
        self.contact_forces = torch.zeros(N, B, 3, **f)
        self.torques = torch.zeros(N, D, **f)
        self.actions = torch.zeros(N, D, **f)
        self.last_actions = torch.zeros(N, D, **f)
        self.last_dof_vel = torch.zeros(N, D, **f)
        self.last_root_vel = torch.zeros(N, 6, **f)
        self.commands = torch.zeros(N, 4, **f)
        self.feet_air_time = torch.zeros(N, len(p.feet_indices), **f)
        self.last_contacts = torch.zeros(N, len(p.feet_indices), dtype=torch.bool, device=device)   # (:547)
        self.episode_length_buf = torch.zeros(N, dtype=torch.long, device=device)
        self.env_origins = torch.zeros(N, 3, **f)
        self.terrain_levels = torch.zeros(N, dtype=torch.long, device=device)
        self.obs_buf = torch.zeros(N, p.num_obs, **f)
        self.rew_buf = torch.zeros(N, **f)
        self.reset_buf = torch.ones(N, dtype=torch.bool, device=device)
        self.time_out_buf = torch.zeros(N, dtype=torch.bool, device=device)
        self.base_lin_vel = torch.zeros(N, 3, **f)
        self.base_ang_vel = torch.zeros(N, 3, **f)
        self.projected_gravity = torch.zeros(N, 3, **f)
        self.measured_heights = 0
        self.episode_sums = {k: torch.zeros(N, **f) for k in p.reward_scales.keys()}
        self.extras: dict = {}
        self.common_step_counter = 0
        self.init_done = True
        self.reset_env_ids = torch.zeros(0, dtype=torch.long, device=device)
        self.resample_env_ids = torch.zeros(0, dtype=torch.long, device=device)
        self.height_indices = None
        self._p = p

    # views the reference creates once (:520-522)
    @property
    def dof_pos(self):
        return self.dof_state.view(self._p.num_envs, self._p.num_dof, 2)[..., 0]

    @property
    def dof_vel(self):
        return self.dof_state.view(self._p.num_envs, self._p.num_dof, 2)[..., 1]

    @property
    def base_quat(self):
        return self.root_states[:, 3:7]

    @staticmethod
    def from_env(env, p: OracleParams, device=None) -> "OracleState":
        """Copy of an env's current buffers (this repo's LeggedRobot or the reference's)."""
        dev = device or env.device
        s = OracleState(p, dev)
        for name in STATE_TENSORS:
            v = getattr(env, name, None)
            if torch.is_tensor(v):
                cur = getattr(s, name)
                setattr(s, name, v.detach().clone().to(dev).view(cur.shape).to(cur.dtype))
        for k in s.episode_sums:
            s.episode_sums[k] = env.episode_sums[k].detach().clone().to(dev)
        mh = getattr(env, "measured_heights", 0)
        s.measured_heights = mh.detach().clone().to(dev) if torch.is_tensor(mh) else 0
        s.common_step_counter = int(env.common_step_counter)
        s.init_done = bool(env.init_done)
        return s


# ----------------------------------------------------------------------------------------------
# control (a2-a4)
def compute_poses(p: OracleParams, actions):
    """envs/base/legged_robot.py:394-397"""
    actions_scaled = actions * p.action_scale
    target_poses = actions_scaled + p.default_dof_pos
    return torch.clip(target_poses, p.dof_pos_limits[:, 0], p.dof_pos_limits[:, 1])


def compute_torques(p: OracleParams, s: OracleState, actions):
    """envs/base/legged_robot.py:370-392"""
    actions_scaled = actions * p.action_scale
    if p.control_type == "P":
        torques = p.p_gains * (actions_scaled + p.default_dof_pos - s.dof_pos) - p.d_gains * s.dof_vel
    elif p.control_type == "V":
        torques = p.p_gains * (actions_scaled - s.dof_vel) - p.d_gains * (s.dof_vel - s.last_dof_vel) / p.sim_dt
    elif p.control_type == "T":
        torques = actions_scaled
    else:
        raise NameError(f"Unknown controller type: {p.control_type}")
    return torch.clip(torques, -p.torque_limits, p.torque_limits)


class ActuatorOracle:
    """Go1 actuator net: history shift + per-leg MLP (envs/go1/go1.py:22-35,50-53,59-66,79-107)."""
    LEG_NUM, LEG_DOF, LEN_HIST = 4, 3, 5
    MODEL_IN_SIZE = 2 * LEG_DOF * LEN_HIST

    def __init__(self, num_envs, state_dict, device="cpu"):
        self.device = device
        sd = {k: v.detach().float().to(device) for k, v in state_dict.items()}
        self.layers = [(sd[f"architecture.{i}.weight"], sd[f"architecture.{i}.bias"]) for i in (0, 2, 4, 6)]
        t = lambda v: torch.tile(torch.tensor(v), (self.LEG_NUM,)).to(device)
        self.pos_err_mean = t([0.00036437, 0.01540757, -0.00972657])
        self.pos_err_std = t([0.11722939, 0.19275887, 0.28700321])
        self.vel_mean = t([-0.00017714, -0.00024455, 0.0005956])
        self.vel_std = t([2.31517027, 3.84613839, 5.52599008])
        self.model_ins = torch.zeros(num_envs, self.MODEL_IN_SIZE * self.LEG_NUM, device=device)
        self.pos_err_buffs = np.zeros((num_envs, 12, self.LEN_HIST))   # float64 on the host, as in go1.py:65-66
        self.vel_buffs = np.zeros((num_envs, 12, self.LEN_HIST))

    def core_model(self, x):
        for i, (w, b) in enumerate(self.layers):
            x = torch.nn.functional.linear(x, w, b)
            if i < 3:
                x = torch.tanh(x)
        return x

    def advance(self, actions, dof_pos, dof_vel):
        H = self.LEN_HIST
        pos_err = actions - dof_pos                                    # raw actions (go1.py:81)
        pos_err_s = (pos_err - self.pos_err_mean) / self.pos_err_std
        vel_s = (dof_vel - self.vel_mean) / self.vel_std
        for i in range(12):                                            # go1.py:87-98
            self.pos_err_buffs[:, i, :] = np.append(np.delete(self.pos_err_buffs[:, i, :], 0, axis=1),
                                                    pos_err_s[:, i].unsqueeze(-1).cpu().numpy(), axis=1)
            self.vel_buffs[:, i, :] = np.append(np.delete(self.vel_buffs[:, i, :], 0, axis=1),
                                                vel_s[:, i].unsqueeze(-1).cpu().numpy(), axis=1)
            self.model_ins[:, 2 * i * H:(2 * i + 1) * H] = torch.from_numpy(self.pos_err_buffs[:, i, :])
            self.model_ins[:, (2 * i + 1) * H:(2 * i + 2) * H] = torch.from_numpy(self.vel_buffs[:, i, :])
        outs = [self.core_model(self.model_ins[:, self.MODEL_IN_SIZE * l:self.MODEL_IN_SIZE * (l + 1)])
                for l in range(self.LEG_NUM)]                          # UniNet.forward, go1.py:31-34
        dVel = torch.cat(outs, 1)
        return dVel * self.vel_std                                     # go1.py:102-105


class SeaLstmOracle:
    """ANYmal series-elastic actuator net: Anymal._init_buffers / _compute_torques / reset_idx
    (envs/anymal_c/anymal.py:56-81) around the TorchScript module LSTMsea of
    resources/actuator_nets/anydrive_v3_lstm.pt, whose forward is
        x0 = x * in_scale; x1, hcn = lstm(x0, hc0); return out_scale * squeeze(linear(x1)), hcn
    with lstm = torch.nn.LSTM(2, 8, num_layers=2, batch_first=True).  The LSTM cell is written out
    (gate rows i, f, g, o; torch.nn.LSTM documentation) so that no step hides in aten::lstm."""
    HID, LAYERS = 8, 2

    def __init__(self, num_envs, num_actions, state_dict, device="cpu"):
        self.N, self.D = num_envs, num_actions
        self.sd = {k: v.detach().float().to(device) for k, v in state_dict.items()}
        M = num_envs * num_actions
        self.sea_input = torch.zeros(M, 1, 2, device=device)                       # anymal.py:65
        self.sea_hidden_state = torch.zeros(self.LAYERS, M, self.HID, device=device)
        self.sea_cell_state = torch.zeros(self.LAYERS, M, self.HID, device=device)
        self.sea_hidden_state_per_env = self.sea_hidden_state.view(self.LAYERS, num_envs, num_actions, self.HID)
        self.sea_cell_state_per_env = self.sea_cell_state.view(self.LAYERS, num_envs, num_actions, self.HID)

    def network(self, x, hc):
        sd = self.sd
        h0, c0 = hc
        inp = (x * sd["in_scale"])[:, 0, :]                                         # one time step
        hn, cn = [], []
        for l in range(self.LAYERS):
            gates = (torch.nn.functional.linear(inp, sd[f"lstm.weight_ih_l{l}"], sd[f"lstm.bias_ih_l{l}"])
                     + torch.nn.functional.linear(h0[l], sd[f"lstm.weight_hh_l{l}"], sd[f"lstm.bias_hh_l{l}"]))
            i, f, g, o = gates.chunk(4, dim=1)
            c = torch.sigmoid(f) * c0[l] + torch.sigmoid(i) * torch.tanh(g)
            h = torch.sigmoid(o) * torch.tanh(c)
            hn.append(h); cn.append(c)
            inp = h
        x1 = inp.unsqueeze(1)
        out = sd["out_scale"] * torch.squeeze(torch.nn.functional.linear(x1, sd["linear.weight"], sd["linear.bias"]))
        return out, (torch.stack(hn), torch.stack(cn))

    def compute_torques(self, actions, action_scale, default_dof_pos, dof_pos, dof_vel):
        """anymal.py:71-78 (use_actuator_network branch); returns the flat [N*D] network output."""
        self.sea_input[:, 0, 0] = (actions * action_scale + default_dof_pos - dof_pos).flatten()
        self.sea_input[:, 0, 1] = dof_vel.flatten()
        torques, (h, c) = self.network(self.sea_input, (self.sea_hidden_state, self.sea_cell_state))
        self.sea_hidden_state[:] = h
        self.sea_cell_state[:] = c
        return torques

    def reset(self, env_ids):
        """anymal.py:58-60"""
        self.sea_hidden_state_per_env[:, env_ids] = 0.
        self.sea_cell_state_per_env[:, env_ids] = 0.


# ----------------------------------------------------------------------------------------------
# post_physics_step and its parts (a5-a18)
def get_heights(p: OracleParams, s: OracleState):
    """envs/base/legged_robot.py:818-854 (+ :802-816 for the points). Also records the clipped indices."""
    N = p.num_envs
    H = p.height_points.shape[0]
    if p.mesh_type == 'plane':
        return torch.zeros(N, H, device=s.root_states.device)
    elif p.mesh_type == 'none':
        raise NameError("Can't measure height with terrain mesh type 'none'")
    height_points = p.height_points.unsqueeze(0).expand(N, H, 3)
    points = quat_apply_yaw(s.base_quat.repeat(1, H), height_points) + (s.root_states[:, :3]).unsqueeze(1)
    points += p.border_size
    points = (points / p.horizontal_scale).long()
    px = points[:, :, 0].view(-1)
    py = points[:, :, 1].view(-1)
    px = torch.clip(px, 0, p.height_samples.shape[0] - 2)
    py = torch.clip(py, 0, p.height_samples.shape[1] - 2)
    s.height_indices = torch.stack([px.view(N, H), py.view(N, H)], dim=-1)
    heights1 = p.height_samples[px, py]
    heights2 = p.height_samples[px + 1, py]
    heights3 = p.height_samples[px, py + 1]
    heights = torch.min(heights1, heights2)
    heights = torch.min(heights, heights3)
    return heights.view(N, -1) * p.vertical_scale


def resample_commands(p: OracleParams, s: OracleState, env_ids, u):
    """envs/base/legged_robot.py:354-368; u = [N,3] per-env uniforms for this site."""
    cr = p.command_ranges
    s.commands[env_ids, 0] = rand_affine(cr["lin_vel_x"][0], cr["lin_vel_x"][1], u[env_ids, 0:1]).squeeze(1)
    s.commands[env_ids, 1] = rand_affine(cr["lin_vel_y"][0], cr["lin_vel_y"][1], u[env_ids, 1:2]).squeeze(1)
    if p.heading_command:
        s.commands[env_ids, 3] = rand_affine(cr["heading"][0], cr["heading"][1], u[env_ids, 2:3]).squeeze(1)
    else:
        s.commands[env_ids, 2] = rand_affine(cr["ang_vel_yaw"][0], cr["ang_vel_yaw"][1], u[env_ids, 2:3]).squeeze(1)
    s.commands[env_ids, :2] *= (torch.norm(s.commands[env_ids, :2], dim=1) > 0.2).unsqueeze(1)


def post_physics_step_callback(p, s, taps):
    """envs/base/legged_robot.py:337-352"""
    env_ids = (s.episode_length_buf % p.resampling_interval == 0).nonzero(as_tuple=False).flatten()
    s.resample_env_ids = env_ids
    resample_commands(p, s, env_ids, taps["cmd_callback"])
    if p.heading_command:
        forward_vec = torch.tensor([1., 0., 0.], device=s.commands.device).repeat((p.num_envs, 1))
        forward = quat_apply(s.base_quat, forward_vec)
        heading = torch.atan2(forward[:, 1], forward[:, 0])
        s.commands[:, 2] = torch.clip(0.5 * wrap_to_pi(s.commands[:, 3] - heading), -1., 1.)
    if p.measure_heights:
        s.measured_heights = get_heights(p, s)
    if p.push_robots and (s.common_step_counter % p.push_interval == 0):
        s.root_states[:, 7:9] = rand_affine(-p.max_push_vel_xy, p.max_push_vel_xy, taps["push"])   # :436-441


def check_termination(p, s):
    """envs/base/legged_robot.py:143-148"""
    s.reset_buf = torch.any(torch.norm(s.contact_forces[:, p.termination_contact_indices, :], dim=-1) > 1., dim=1)
    s.time_out_buf = s.episode_length_buf > p.max_episode_length
    s.reset_buf |= s.time_out_buf


def reward_terms(p: OracleParams, s: OracleState):
    """The _reward_* functions (envs/base/legged_robot.py:857-966, envs/cassie/cassie.py:43-46)."""
    cf, feet = s.contact_forces, p.feet_indices
    d = dict
    def feet_air_time():
        contact = cf[:, feet, 2] > 1.
        if p.filtered_feet_contact:
            # the variant the reference keeps commented out (:928-939): PhysX contact reporting is unreliable on
            # meshes, so a foot counts as in contact if it is now or was at the previous step.  No executable
            # reference exists for it: PARITY UNPINNED for this option (restated from the commented lines)
            now = contact
            contact = torch.logical_or(now, s.last_contacts)
            s.last_contacts = now
        first_contact = (s.feet_air_time > 0.) * contact
        s.feet_air_time += p.dt
        rew = torch.sum((s.feet_air_time - 0.5) * first_contact, dim=1)
        rew *= torch.norm(s.commands[:, :2], dim=1) > 0.1
        s.feet_air_time *= ~contact
        return rew

    def dof_pos_limits():
        out = -(s.dof_pos - p.dof_pos_limits[:, 0]).clip(max=0.)
        out += (s.dof_pos - p.dof_pos_limits[:, 1]).clip(min=0.)
        return torch.sum(out, dim=1)

    def base_height():
        base_height = torch.mean(s.root_states[:, 2].unsqueeze(1) - s.measured_heights, dim=1)
        return torch.square(base_height - p.base_height_target)

    def no_fly():
        contacts = cf[:, feet, 2] > 0.1
        return 1. * (torch.sum(1. * contacts, dim=1) == 1)

    return d(
        lin_vel_z=lambda: torch.square(s.base_lin_vel[:, 2]),
        ang_vel_xy=lambda: torch.sum(torch.square(s.base_ang_vel[:, :2]), dim=1),
        orientation=lambda: torch.sum(torch.square(s.projected_gravity[:, :2]), dim=1),
        base_height=base_height,
        torques=lambda: torch.sum(torch.square(s.torques), dim=1),
        energy=lambda: torch.sum(torch.square(s.torques * s.dof_vel), dim=1),
        dof_vel=lambda: torch.sum(torch.square(s.dof_vel), dim=1),
        dof_acc=lambda: torch.sum(torch.square((s.last_dof_vel - s.dof_vel) / p.dt), dim=1),
        action_rate=lambda: torch.sum(torch.square(s.last_actions - s.actions), dim=1),
        collision=lambda: torch.sum(1. * (torch.norm(cf[:, p.penalised_contact_indices, :], dim=-1) > 0.1), dim=1),
        termination=lambda: s.reset_buf * ~s.time_out_buf,
        dof_pos_limits=dof_pos_limits,
        dof_vel_limits=lambda: torch.sum((torch.abs(s.dof_vel) - p.dof_vel_limits * p.soft_dof_vel_limit).clip(min=0., max=1.), dim=1),
        torque_limits=lambda: torch.sum((torch.abs(s.torques) - p.torque_limits * p.soft_torque_limit).clip(min=0.), dim=1),
        tracking_lin_vel=lambda: torch.exp(-torch.sum(torch.square(s.commands[:, :2] - s.base_lin_vel[:, :2]), dim=1) / p.tracking_sigma),
        tracking_ang_vel=lambda: torch.exp(-torch.square(s.commands[:, 2] - s.base_ang_vel[:, 2]) / p.tracking_sigma),
        feet_air_time=feet_air_time,
        stumble=lambda: torch.any(torch.norm(cf[:, feet, :2], dim=2) > 5 * torch.abs(cf[:, feet, 2]), dim=1),
        stand_still=lambda: torch.sum(torch.abs(s.dof_pos - p.default_dof_pos), dim=1) * (torch.norm(s.commands[:, :2], dim=1) < 0.1),
        feet_contact_forces=lambda: torch.sum((torch.norm(cf[:, feet, :], dim=-1) - p.max_contact_force).clip(min=0.), dim=1),
        hip_motion=lambda: torch.sum(torch.abs(s.dof_pos[:, [0, 3, 6, 9]] - p.default_dof_pos[:, [0, 3, 6, 9]]), dim=1),
        no_fly=no_fly,
    )


def compute_reward(p, s):
    """envs/base/legged_robot.py:195-212"""
    terms = reward_terms(p, s)
    for name, fn in (p.extra_reward_terms or {}).items():
        terms[name] = (lambda f=fn: f(p, s))
    s.rew_buf[:] = 0.
    for name in p.reward_names:
        rew = terms[name]() * p.reward_scales[name]
        s.rew_buf += rew
        s.episode_sums[name] += rew
    if p.only_positive_rewards:
        s.rew_buf[:] = torch.clip(s.rew_buf[:], min=0.)
    if "termination" in p.reward_scales:
        rew = terms["termination"]() * p.reward_scales["termination"]
        s.rew_buf += rew
        s.episode_sums["termination"] += rew


def update_terrain_curriculum(p, s, env_ids, taps):
    """envs/base/legged_robot.py:443-463"""
    if not s.init_done:
        return
    distance = torch.norm(s.root_states[env_ids, :2] - s.env_origins[env_ids, :2], dim=1)
    move_up = distance > p.env_length / 2
    move_down = (distance < torch.norm(s.commands[env_ids, :2], dim=1) * p.max_episode_length_s * 0.5) * ~move_up
    s.terrain_levels[env_ids] += 1 * move_up - 1 * move_down
    s.terrain_levels[env_ids] = torch.where(s.terrain_levels[env_ids] >= p.max_terrain_level,
                                            taps["level"][env_ids],
                                            torch.clip(s.terrain_levels[env_ids], 0))
    s.env_origins[env_ids] = p.terrain_origins[s.terrain_levels[env_ids], p.terrain_types[env_ids]]


def update_command_curriculum(p, s, env_ids):
    """envs/base/legged_robot.py:465-474"""
    if torch.mean(s.episode_sums["tracking_lin_vel"][env_ids]) / p.max_episode_length > 0.8 * p.reward_scales["tracking_lin_vel"]:
        p.command_ranges["lin_vel_x"][0] = np.clip(p.command_ranges["lin_vel_x"][0] - 0.5, -p.max_curriculum, 0.)
        p.command_ranges["lin_vel_x"][1] = np.clip(p.command_ranges["lin_vel_x"][1] + 0.5, 0., p.max_curriculum)


def reset_idx(p, s, env_ids, taps):
    """envs/base/legged_robot.py:150-193 with _reset_dofs :399-408 and _reset_root_states :414-430"""
    s.reset_env_ids = env_ids
    if len(env_ids) == 0:
        return
    if p.terrain_curriculum:
        update_terrain_curriculum(p, s, env_ids, taps)
    if p.commands_curriculum and (s.common_step_counter % p.max_episode_length == 0):
        update_command_curriculum(p, s, env_ids)
    # _reset_dofs
    s.dof_pos[env_ids] = p.default_dof_pos * rand_affine(0.5, 1.5, taps["dof"][env_ids])
    s.dof_vel[env_ids] = 0.
    # _reset_root_states
    s.root_states[env_ids] = p.base_init_state
    s.root_states[env_ids, :3] += s.env_origins[env_ids]
    if p.custom_origins:
        s.root_states[env_ids, :2] += rand_affine(-1., 1., taps["root_xy"][env_ids])
    s.root_states[env_ids, 7:13] = rand_affine(-0.5, 0.5, taps["root_vel"][env_ids])
    resample_commands(p, s, env_ids, taps["cmd_reset"])
    s.last_actions[env_ids] = 0.
    s.last_dof_vel[env_ids] = 0.
    s.feet_air_time[env_ids] = 0.
    s.episode_length_buf[env_ids] = 0
    s.reset_buf[env_ids] = 1
    s.extras["episode"] = {}
    for key in s.episode_sums.keys():
        s.extras["episode"]['rew_' + key] = torch.mean(s.episode_sums[key][env_ids]) / p.max_episode_length_s
        s.episode_sums[key][env_ids] = 0.
    if p.terrain_curriculum:
        s.extras["episode"]["terrain_level"] = torch.mean(s.terrain_levels.float())
    if p.commands_curriculum:
        s.extras["episode"]["max_command_x"] = p.command_ranges["lin_vel_x"][1]
    if p.send_timeouts:
        s.extras["time_outs"] = s.time_out_buf


def compute_observations(p, s, taps):
    """envs/base/legged_robot.py:214-231"""
    sc = p.obs_scales
    s.obs_buf = torch.cat((s.base_lin_vel * sc["lin_vel"],
                           s.base_ang_vel * sc["ang_vel"],
                           s.projected_gravity,
                           s.commands[:, :3] * p.commands_scale,
                           (s.dof_pos - p.default_dof_pos) * sc["dof_pos"],
                           s.dof_vel * sc["dof_vel"],
                           s.actions), dim=-1)
    if p.measure_heights:
        heights = torch.clip(s.root_states[:, 2].unsqueeze(1) - 0.5 - s.measured_heights, -1, 1.) * sc["height_measurements"]
        s.obs_buf = torch.cat((s.obs_buf, heights), dim=-1)
    if p.add_noise:
        s.obs_buf += (2 * taps["noise"] - 1) * p.noise_scale_vec


def post_physics_step(p: OracleParams, s: OracleState, taps):
    """envs/base/legged_robot.py:109-141"""
    s.episode_length_buf += 1
    s.common_step_counter += 1
    gravity_vec = torch.tensor([0., 0., -1.], device=s.root_states.device).repeat((p.num_envs, 1))
    s.base_lin_vel = quat_rotate_inverse(s.base_quat, s.root_states[:, 7:10])
    s.base_ang_vel = quat_rotate_inverse(s.base_quat, s.root_states[:, 10:13])
    s.projected_gravity = quat_rotate_inverse(s.base_quat, gravity_vec)
    post_physics_step_callback(p, s, taps)
    check_termination(p, s)
    compute_reward(p, s)
    env_ids = s.reset_buf.nonzero(as_tuple=False).flatten()
    reset_idx(p, s, env_ids, taps)
    compute_observations(p, s, taps)
    s.last_actions[:] = s.actions[:]
    s.last_dof_vel[:] = s.dof_vel[:]
    s.last_root_vel[:] = s.root_states[:, 7:13]


def step(p: OracleParams, s: OracleState, actions, taps, frame_loader=None):
    """envs/base/legged_robot.py:79-107 without the simulator: `frame_loader(s, substep)` stands in for
    simulate + refresh.  Returns the target poses of the last substep."""
    s.actions = torch.clip(actions, -p.clip_actions, p.clip_actions).to(s.root_states.device)
    target = None
    for k in range(p.decimation):
        target = compute_poses(p, s.actions)
        if frame_loader is not None:
            frame_loader(s, k)
    post_physics_step(p, s, taps)
    s.obs_buf = torch.clip(s.obs_buf, -p.clip_observations, p.clip_observations)
    return target
