"""LeggedRobot -- the vectorised locomotion env, B200-native.

Same class, constructor, ``step()/reset()/get_observations()`` contract, public tensors and
``cfg.rewards.scales -> _reward_<name>`` registration as the reference
(``/root/reference/legged_gym/envs/base/legged_robot.py:51-974``).  What differs is where the
arithmetic runs: the bodies of ``step()`` / ``post_physics_step()`` are calls into liblgk.so
(``include/lgk.h``) -- one fused sm_100a kernel per ``post_physics_step`` plus one small control
kernel per simulator substep -- instead of ~350 torch kernel launches and three host syncs.

There is no CPU or torch fallback: an env can be *constructed* on ``device='cpu'`` (host logic,
tables) but ``step()`` needs a CUDA device and the built library.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional

import numpy as np
import torch

from legged_gym_b200 import LEGGED_GYM_ROOT_DIR, _lgk
from legged_gym_b200.envs.base.base_task import BaseTask, HAVE_ISAACGYM, gymapi
from legged_gym_b200.utils.helpers import class_to_dict
from legged_gym_b200.utils.terrain import Terrain

from legged_gym_b200.sim.backend import gymtorch

# hooks whose bodies live inside the fused kernel; a subclass overriding one of them would silently
# be ignored, so construction fails instead
_FUSED_HOOKS = ("compute_reward",
                "_resample_commands", "_get_heights", "_push_robots", "_reset_dofs", "_reset_root_states",
                "_update_terrain_curriculum", "post_physics_step")


class LeggedRobot(BaseTask):
    def __init__(self, cfg, sim_params, physics_engine, sim_device, headless):
        self.cfg = cfg
        self.sim_params = sim_params
        self.height_samples = None
        self.debug_viz = False
        self.init_done = False
        self._episode_length_buf = None
        self._parse_cfg(self.cfg)
        super().__init__(self.cfg, sim_params, physics_engine, sim_device, headless)
        self._init_buffers()
        self._prepare_reward_function()
        self._check_hooks()
        self._init_observation_override()
        self._lib = None
        self._ranges_cache = None
        self._ranges_split = False
        self._consts = None
        self._bufs = None
        self.init_done = True

    # ------------------------------------------------------------------ rsl_rl rebinding
    # OnPolicyRunner.learn(init_at_random_ep_len=True) *re-binds* env.episode_length_buf; the kernels
    # hold the address of the original buffer, so assignment copies into it instead.
    @property
    def episode_length_buf(self):
        return self._episode_length_buf

    @episode_length_buf.setter
    def episode_length_buf(self, value):
        if self._episode_length_buf is None:
            self._episode_length_buf = value
        else:
            self._episode_length_buf.copy_(value)

    # ------------------------------------------------------------------ step
    def step(self, actions):
        """clip actions, `decimation` x (control -> simulate), post_physics_step (legged_robot.py:79-107)."""
        lib = self._kernels()
        actions = actions.to(self.device, dtype=torch.float).contiguous()
        st = _lgk.current_stream(self.device)
        # the clip and the position targets of the clipped actions in one launch; _compute_poses() (and a
        # subclass override that ends in super()._compute_poses(), like Go1's) finds the targets ready
        _lgk.check("lgk_clip_actions_poses", lib.lgk_clip_actions_poses(
            C.byref(self._consts), _lgk.ptr(actions), _lgk.ptr(self.actions), _lgk.ptr(self._target_poses_buf), st))
        self._poses_fresh = True
        self.render()
        for _ in range(self.cfg.control.decimation):
            # position control, as in this fork of the reference (:93-95)
            self.target_poses = self._compute_poses(self.actions).view(self.target_poses.shape)
            self.gym.set_dof_position_target_tensor(self.sim, gymtorch.unwrap_tensor(self.target_poses))
            self.gym.simulate(self.sim)
            if self.device == 'cpu':
                self.gym.fetch_results(self.sim, True)
            self.gym.refresh_dof_state_tensor(self.sim)
        self.post_physics_step()
        # observations are clipped to +-clip_observations inside the kernel (:103-104)
        return self.obs_buf, self.privileged_obs_buf, self.rew_buf, self.reset_buf, self.extras

    def post_physics_step(self):
        """Everything the reference does here (:109-141) in one kernel launch."""
        lib = self._kernels()
        self.gym.refresh_actor_root_state_tensor(self.sim)
        self.gym.refresh_net_contact_force_tensor(self.sim)
        self.gym.refresh_dof_force_tensor(self.sim)
        self.common_step_counter += 1
        self._sync_host_scalars()
        self._ensure_dev_counter(self.common_step_counter - 1)
        if self._callback_override:
            # a subclass's _post_physics_step_callback(): the base part (:337-349: command resample, heading
            # command, height scan) is done by lgk_pre_reward_state, the override then works on self.commands /
            # self.measured_heights in place; the push (:351) stays with the step kernel
            self._ensure_pre_reward_state(lib)
            bound = self.commands
            self._post_physics_step_callback()
            if self.commands is not bound:
                bound.copy_(self.commands.view_as(bound))
                self.commands = bound
            self._consts.flags |= _lgk.F_CALLBACK_GIVEN
        if self._termination_override:
            # a subclass's check_termination(): the base decision (:143-148) is in reset_buf / time_out_buf after
            # lgk_pre_reward_state, the override adds its own conditions, the kernel acts on the result
            self._ensure_pre_reward_state(lib)
            bound_reset, bound_time_out = self.reset_buf, self.time_out_buf
            self.check_termination()
            if self.reset_buf is not bound_reset:           # the override re-bound the attribute (as :146 does)
                bound_reset.copy_(self.reset_buf.to(bound_reset.dtype).view_as(bound_reset))
                self.reset_buf = bound_reset
            if self.time_out_buf is not bound_time_out:
                bound_time_out.copy_(self.time_out_buf.to(bound_time_out.dtype).view_as(bound_time_out))
                self.time_out_buf = bound_time_out
            self._consts.flags |= _lgk.F_RESET_GIVEN
        if self.cfg.commands.curriculum and (self.common_step_counter % self.max_episode_length == 0):
            self._command_curriculum_prepass(lib)
        if self._external_terms:
            self._evaluate_external_terms(lib)
        _lgk.check("lgk_post_physics_step", lib.lgk_post_physics_step(
            C.byref(self._consts), C.byref(self._bufs), _lgk.current_stream(self.device)))
        self._consts.flags &= ~(_lgk.F_EP_INCREMENTED | _lgk.F_RESET_GIVEN | _lgk.F_CALLBACK_GIVEN)
        if self._ranges_split:
            self._sync_host_scalars()          # both halves equal again from the next step on
        self._dev_step_counter = self.common_step_counter
        self._after_step_sim_writes()
        if self._obs_override:
            self._run_observation_override()

    def _evaluate_external_terms(self, lib):
        """Reward terms a subclass defines in Python (`_reward_<name>`, found like the reference does,
        :588-593) that the kernel has no code for.  `lgk_pre_reward_state` first brings the buffers such a
        term may read to the state the reference has inside compute_reward() -- episode_length_buf,
        base_lin_vel / base_ang_vel / projected_gravity, commands after the callback, measured_heights,
        reset_buf / time_out_buf -- then each term is evaluated and its values enter the fused kernel as an
        LGK_REW_EXTERNAL row: scaled, summed in the term's alphabetical slot, accumulated in episode_sums.
        Not reproduced for such terms: feet_air_time as advanced by a built-in term earlier in the order,
        and the pushed root velocities on push steps (both are updated by the fused kernel afterwards)."""
        self._ensure_pre_reward_state(lib)
        i = 0
        for name, fn in zip(self.reward_names, self.reward_functions):
            if fn is not None:
                self._ext_rewards[i].copy_(fn().to(torch.float).reshape(self.num_envs))
                i += 1

    # ------------------------------------------------------------------ control
    def _compute_poses(self, actions):
        """target = clip(a * action_scale + default, soft limits) (:394-397).

        The targets depend on the actions only, so within one step() the kernel runs on the first
        substep and later substeps reuse the buffer."""
        if not (self._poses_fresh and actions is self.actions):
            lib = self._kernels()
            _lgk.check("lgk_compute_poses", lib.lgk_compute_poses(
                C.byref(self._consts), _lgk.ptr(actions), _lgk.ptr(self._target_poses_buf),
                _lgk.current_stream(self.device)))
            self._poses_fresh = actions is self.actions
        return self._target_poses_buf

    def _compute_torques(self, actions):
        """PD torques for control_type P / V / T, clipped to the torque limits (:370-392)."""
        control_type = self.cfg.control.control_type
        if control_type not in _lgk.CTRL:
            raise NameError(f"Unknown controller type: {control_type}")
        lib = self._kernels()
        self._consts.control_type = _lgk.CTRL[control_type]
        _lgk.check("lgk_compute_torques", lib.lgk_compute_torques(
            C.byref(self._consts), _lgk.ptr(actions), _lgk.ptr(self.dof_state), _lgk.ptr(self.last_dof_vel),
            _lgk.ptr(self._torques_out), _lgk.current_stream(self.device)))
        return self._torques_out

    # ------------------------------------------------------------------ resets
    def reset_idx(self, env_ids):
        """Reset the given envs (:150-193). The per-step resets are done by the fused kernel; this
        entry serves BaseTask.reset() and manual resets."""
        if len(env_ids) == 0:
            return
        lib = self._kernels()
        if self.cfg.commands.curriculum and (self.common_step_counter % self.max_episode_length == 0):
            self.update_command_curriculum(env_ids)
        self._sync_host_scalars()
        self._ensure_dev_counter(self.common_step_counter)
        n = int(len(env_ids))
        ids = env_ids.to(device=self.device, dtype=torch.long).contiguous()
        _lgk.check("lgk_reset_idx", lib.lgk_reset_idx(
            C.byref(self._consts), C.byref(self._bufs), _lgk.ptr(ids), n, _lgk.current_stream(self.device)))
        ids32 = ids.to(torch.int32)
        self.gym.set_dof_state_tensor_indexed(self.sim, gymtorch.unwrap_tensor(self.dof_state),
                                              gymtorch.unwrap_tensor(ids32), n)
        self.gym.set_actor_root_state_tensor_indexed(self.sim, gymtorch.unwrap_tensor(self.root_states),
                                                     gymtorch.unwrap_tensor(ids32), n)
        self._publish_extras()

    def update_command_curriculum(self, env_ids):
        """Widen the lin_vel_x command range when tracking is good (:465-474). Host-side, rare."""
        mean = torch.mean(self.episode_sums["tracking_lin_vel"][env_ids]) / self.max_episode_length
        if mean > 0.8 * self.reward_scales["tracking_lin_vel"]:
            r = self.command_ranges["lin_vel_x"]
            r[0] = np.clip(r[0] - 0.5, -self.cfg.commands.max_curriculum, 0.)
            r[1] = np.clip(r[1] + 0.5, 0., self.cfg.commands.max_curriculum)

    def _ensure_pre_reward_state(self, lib):
        """Launch lgk_pre_reward_state once per step (python reward terms and the command curriculum's
        gated step both need the compute_reward-time state before the fused launch)."""
        if not (self._consts.flags & _lgk.F_EP_INCREMENTED):
            _lgk.check("lgk_pre_reward_state", lib.lgk_pre_reward_state(
                C.byref(self._consts), C.byref(self._bufs), _lgk.current_stream(self.device)))
            self._consts.flags |= _lgk.F_EP_INCREMENTED

    def _command_curriculum_prepass(self, lib):
        """The reference updates command_ranges inside reset_idx -- after compute_reward, before the reset
        envs resample (:166-173) -- from the episode sums that already include this step's reward.  The
        fused kernel cannot stop mid-launch for a host decision, so on the one step in max_episode_length
        where the gate is open the decision is formed here first, from the same quantities: the reset set
        and the post-callback commands / body-frame velocity that lgk_pre_reward_state leaves, and this
        step's tracking_lin_vel term added to the sums with the reference's own operations (:918-921, :207)."""
        self._ensure_pre_reward_state(lib)
        env_ids = self.reset_buf.nonzero(as_tuple=False).flatten()
        sharded = self._shard_world() > 1
        if len(env_ids) == 0 and not sharded:
            return
        sums = self.episode_sums["tracking_lin_vel"]          # KeyError without the term, like the reference (:472)
        row = self.reward_names.index("tracking_lin_vel")
        fn = self.reward_functions[row]
        if fn is not None:
            rew = fn()
        else:
            lin_vel_error = torch.sum(torch.square(self.commands[:, :2] - self.base_lin_vel[:, :2]), dim=1)
            rew = torch.exp(-lin_vel_error / self.cfg.rewards.tracking_sigma)
        sums = sums + rew * self.reward_scales["tracking_lin_vel"]
        mean = self._mean_over_all_shards(sums[env_ids]) if sharded else torch.mean(sums[env_ids])
        if mean is not None and mean / self.max_episode_length > 0.8 * self.reward_scales["tracking_lin_vel"]:
            r = self.command_ranges["lin_vel_x"]
            r[0] = np.clip(r[0] - 0.5, -self.cfg.commands.max_curriculum, 0.)
            r[1] = np.clip(r[1] + 0.5, 0., self.cfg.commands.max_curriculum)
            self._sync_host_scalars(reset_half_only=True)   # the callback's resample still sees the old range

    def _shard_world(self) -> int:
        import torch.distributed as dist
        world = int(getattr(self.cfg.env, "shard_world", 1))
        return world if (world > 1 and dist.is_available() and dist.is_initialized()) else 1

    @staticmethod
    def _mean_over_all_shards(values):
        """Mean of `values` over the envs of every rank (one all-reduce of [sum, count]): the command
        curriculum's gate (:472) then decides on every GPU what a single process holding all the envs would
        decide, and command_ranges stay identical across ranks.  Returns None when no rank has a value (the
        reference does not call update_command_curriculum on an empty reset set, :160)."""
        import torch.distributed as dist
        t = torch.stack([values.sum().float(), torch.tensor(float(values.numel()), device=values.device)])
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return None if float(t[1]) == 0.0 else t[0] / t[1]

    def _after_step_sim_writes(self):
        """Hand the state the kernel rewrote back to the simulator (:411,432,441)."""
        push = self.cfg.domain_rand.push_robots and (self.common_step_counter % self.cfg.domain_rand.push_interval == 0)
        if HAVE_ISAACGYM:
            if push:
                self.gym.set_actor_root_state_tensor(self.sim, gymtorch.unwrap_tensor(self.root_states))
            n = int(self.step_counts[0].item())   # the simulator API wants a host count
            if n > 0:
                ids32 = self.reset_ids[:n]
                self.gym.set_dof_state_tensor_indexed(self.sim, gymtorch.unwrap_tensor(self.dof_state),
                                                      gymtorch.unwrap_tensor(ids32), n)
                self.gym.set_actor_root_state_tensor_indexed(self.sim, gymtorch.unwrap_tensor(self.root_states),
                                                             gymtorch.unwrap_tensor(ids32), n)
        else:
            # headless backend: nothing consumes the writes; keep the step free of host syncs
            if push:
                self.gym.set_actor_root_state_tensor(self.sim, self.root_states)
            self.gym.set_dof_state_tensor_indexed(self.sim, self.dof_state, self.reset_ids, self.step_counts)
            self.gym.set_actor_root_state_tensor_indexed(self.sim, self.root_states, self.reset_ids, self.step_counts)
        self._after_fused_resets()
        self._publish_extras()

    def _after_fused_resets(self):
        """Hook for subclasses whose reset_idx() override clears extra per-env state (Anymal's LSTM state,
        anymal.py:56-60).  The per-step resets run inside the fused kernel, which leaves the ascending id
        list in self.reset_ids and its length in self.step_counts[0] (both on the device)."""

    def _publish_extras(self):
        """extras["episode"] / extras["time_outs"] as device views (:182-193); the kernel rewrites the
        underlying values only on steps with at least one reset, like the reference."""
        if "episode" not in self.extras:
            ep = {}
            for name, row in self._sum_row.items():
                ep['rew_' + name] = self._extras_buf[row]
            if self.cfg.terrain.curriculum:
                ep["terrain_level"] = self._extras_buf[self._num_sum_rows]
            self.extras["episode"] = ep
        if self.cfg.commands.curriculum:
            self.extras["episode"]["max_command_x"] = self.command_ranges["lin_vel_x"][1]
        if self.cfg.env.send_timeouts:
            self.extras["time_outs"] = self.time_out_buf

    # ------------------------------------------------------------------ sim creation
    def create_sim(self):
        self.up_axis_idx = 2
        self.sim = self.gym.create_sim(self.sim_device_id, self.graphics_device_id, self.physics_engine, self.sim_params)
        mesh_type = self.cfg.terrain.mesh_type
        if mesh_type in ['heightfield', 'trimesh']:
            self.terrain = Terrain(self.cfg.terrain, self.num_envs, device=self.device)
        if mesh_type == 'plane':
            self._create_ground_plane()
        elif mesh_type == 'heightfield':
            self._create_heightfield()
        elif mesh_type == 'trimesh':
            self._create_trimesh()
        elif mesh_type is not None:
            raise ValueError("Terrain mesh type not recognised. Allowed types are [None, plane, heightfield, trimesh]")
        self._create_envs()

    def set_camera(self, position, lookat):
        raise RuntimeError("no viewer: headless only")

    # ---- callbacks kept from the reference (one-time setup) --------------------------------
    def _process_rigid_shape_props(self, props, env_id):
        if self.cfg.domain_rand.randomize_friction:
            if env_id == 0:
                lo, hi = self.cfg.domain_rand.friction_range
                num_buckets = 64
                bucket_ids = torch.randint(0, num_buckets, (self.num_envs, 1))
                buckets = (hi - lo) * torch.rand(num_buckets, 1) + lo
                self.friction_coeffs = buckets[bucket_ids]
            for s in range(len(props)):
                props[s].friction = self.friction_coeffs[env_id]
        return props

    def _process_dof_props(self, props, env_id):
        """Stores URDF limits and derives the soft position limits (:284-310), in fp32 like the reference."""
        if env_id == 0:
            kw = dict(dtype=torch.float, device=self.device, requires_grad=False)
            self.dof_pos_limits = torch.zeros(self.num_dof, 2, **kw)
            self.dof_vel_limits = torch.zeros(self.num_dof, **kw)
            self.torque_limits = torch.zeros(self.num_dof, **kw)
            soft = self.cfg.rewards.soft_dof_pos_limit
            for i in range(len(props)):
                self.dof_pos_limits[i, 0] = props["lower"][i].item()
                self.dof_pos_limits[i, 1] = props["upper"][i].item()
                self.dof_vel_limits[i] = props["velocity"][i].item()
                self.torque_limits[i] = props["effort"][i].item()
                m = (self.dof_pos_limits[i, 0] + self.dof_pos_limits[i, 1]) / 2
                r = self.dof_pos_limits[i, 1] - self.dof_pos_limits[i, 0]
                self.dof_pos_limits[i, 0] = m - 0.5 * r * soft
                self.dof_pos_limits[i, 1] = m + 0.5 * r * soft
        return props

    def _process_rigid_body_props(self, props, env_id):
        dr = self.cfg.domain_rand
        for i, p in enumerate(props):
            if i == 0:
                if dr.randomize_base_mass:
                    p.mass += np.random.uniform(dr.added_mass_range[0], dr.added_mass_range[1])
            elif dr.randomize_limb_mass:
                p.mass *= (1 + np.random.uniform(dr.added_limb_percentage[0], dr.added_limb_percentage[1]))
        return props

    def _get_noise_scale_vec(self, cfg):
        """Per-column noise amplitude (:477-500); the slices assume 12 joints like the reference."""
        noise_vec = torch.zeros_like(self.obs_buf[0])
        self.add_noise = self.cfg.noise.add_noise
        ns = self.cfg.noise.noise_scales
        lvl = self.cfg.noise.noise_level
        noise_vec[:3] = ns.lin_vel * lvl * self.obs_scales.lin_vel
        noise_vec[3:6] = ns.ang_vel * lvl * self.obs_scales.ang_vel
        noise_vec[6:9] = ns.gravity * lvl
        noise_vec[9:12] = 0.
        noise_vec[12:24] = ns.dof_pos * lvl * self.obs_scales.dof_pos
        noise_vec[24:36] = ns.dof_vel * lvl * self.obs_scales.dof_vel
        noise_vec[36:48] = 0.
        if self.cfg.terrain.measure_heights:
            noise_vec[48:235] = ns.height_measurements * lvl * self.obs_scales.height_measurements
        return noise_vec

    # ------------------------------------------------------------------ buffers
    def _init_buffers(self):
        """Wrap the simulator tensors and allocate the env state (:503-572)."""
        gym, sim = self.gym, self.sim
        actor_root_state = gym.acquire_actor_root_state_tensor(sim)
        dof_state_tensor = gym.acquire_dof_state_tensor(sim)
        net_contact_forces = gym.acquire_net_contact_force_tensor(sim)
        torques = gym.acquire_dof_force_tensor(sim)
        gym.refresh_dof_state_tensor(sim)
        gym.refresh_actor_root_state_tensor(sim)
        gym.refresh_net_contact_force_tensor(sim)
        gym.refresh_dof_force_tensor(sim)

        N, D = self.num_envs, self.num_dof
        self.root_states = gymtorch.wrap_tensor(actor_root_state)
        self.dof_state = gymtorch.wrap_tensor(dof_state_tensor)
        self.dof_pos = self.dof_state.view(N, D, 2)[..., 0]
        self.dof_vel = self.dof_state.view(N, D, 2)[..., 1]
        self.base_quat = self.root_states[:, 3:7]
        self.contact_forces = gymtorch.wrap_tensor(net_contact_forces).view(N, -1, 3)

        kw = dict(dtype=torch.float, device=self.device, requires_grad=False)
        self.common_step_counter = 0
        self._dev_step_counter = 0
        self.extras = {}
        self.noise_scale_vec = self._get_noise_scale_vec(self.cfg)
        self.gravity_vec = torch.tensor([0., 0., -1.], **kw).repeat((N, 1))
        self.forward_vec = torch.tensor([1., 0., 0.], **kw).repeat((N, 1))
        self._target_poses_buf = torch.zeros(N, self.num_actions, **kw)
        self.target_poses = self._target_poses_buf
        self._torques_out = torch.zeros(N, self.num_actions, **kw)
        self.torques = gymtorch.wrap_tensor(torques).view(N, self.num_actions)   # simulator-reported (:536)
        self.p_gains = torch.zeros(self.num_actions, **kw)
        self.d_gains = torch.zeros(self.num_actions, **kw)
        self.actions = torch.zeros(N, self.num_actions, **kw)
        self.last_actions = torch.zeros(N, self.num_actions, **kw)
        self.last_dof_vel = torch.zeros(N, D, **kw)
        self.last_root_vel = torch.zeros(N, 6, **kw)
        self.commands = torch.zeros(N, self.cfg.commands.num_commands, **kw)
        self.commands_scale = torch.tensor([self.obs_scales.lin_vel, self.obs_scales.lin_vel, self.obs_scales.ang_vel],
                                           device=self.device, requires_grad=False)
        self.feet_air_time = torch.zeros(N, self.feet_indices.shape[0], **kw)
        self.last_contacts = torch.zeros(N, len(self.feet_indices), dtype=torch.bool, device=self.device)
        self.base_lin_vel = torch.zeros(N, 3, **kw)
        self.base_ang_vel = torch.zeros(N, 3, **kw)
        self.projected_gravity = torch.zeros(N, 3, **kw)
        self.projected_gravity[:, 2] = -1.
        if self.cfg.terrain.measure_heights:
            self.height_points = self._init_height_points()
        # the reference keeps the python int 0 here until the first scan (:553); a [N,H] buffer is
        # materialised only when cfg.terrain.materialize_heights (default True) asks for it
        self.measured_heights = 0
        if self.cfg.terrain.measure_heights and getattr(self.cfg.terrain, "materialize_heights", True):
            self.measured_heights = torch.zeros(N, self.num_height_points, **kw)

        self.default_dof_pos = torch.zeros(D, **kw)
        for i in range(self.num_dofs):
            name = self.dof_names[i]
            self.default_dof_pos[i] = self.cfg.init_state.default_joint_angles[name]
            found = False
            for dof_name in self.cfg.control.stiffness.keys():
                if dof_name in name:
                    self.p_gains[i] = self.cfg.control.stiffness[dof_name]
                    self.d_gains[i] = self.cfg.control.damping[dof_name]
                    found = True
            if not found:
                self.p_gains[i] = 0.
                self.d_gains[i] = 0.
                if self.cfg.control.control_type in ["P", "V"]:
                    print(f"PD gain of joint {name} were not defined, setting them to zero")
        self.default_dof_pos = self.default_dof_pos.unsqueeze(0)

        # outputs of the compaction / statistics
        self.reset_ids = torch.zeros(N, dtype=torch.int32, device=self.device)
        self.resample_ids = torch.zeros(N, dtype=torch.int32, device=self.device)
        self.step_counts = torch.zeros(2, dtype=torch.int32, device=self.device)
        self._poses_fresh = False

    def _prepare_reward_function(self):
        """Non-zero scales * dt, bound to reward terms in alphabetical order (:574-598)."""
        for key in list(self.reward_scales.keys()):
            if self.reward_scales[key] == 0:
                self.reward_scales.pop(key)
            else:
                self.reward_scales[key] *= self.dt
        self.reward_names: List[str] = []
        self.reward_functions = []
        self._reward_term_ids: List[int] = []
        self._external_terms: List[str] = []
        for name in self.reward_scales.keys():
            if name == "termination":
                continue
            user_fn = getattr(self, '_reward_' + name, None)
            if user_fn is not None:
                # a python _reward_<name> defined by a subclass: evaluated on the host side and summed
                # by the kernel in its alphabetical slot
                self._reward_term_ids.append(_lgk.REW_EXTERNAL)
                self._external_terms.append(name)
                self.reward_functions.append(user_fn)
            elif name in _lgk.REWARD_ID:
                self._reward_term_ids.append(_lgk.REWARD_ID[name])
                self.reward_functions.append(None)   # evaluated inside the fused kernel
            else:
                # same failure as the reference's getattr(self, '_reward_' + name) (:593)
                raise AttributeError(f"'{type(self).__name__}' object has no attribute '_reward_{name}'")
            self.reward_names.append(name)
        rows = list(self.reward_names) + (["termination"] if "termination" in self.reward_scales else [])
        self._sum_row: Dict[str, int] = {n: i for i, n in enumerate(rows)}
        self._num_sum_rows = len(rows)
        self._episode_sums_buf = torch.zeros(max(self._num_sum_rows, 1), self.num_envs, dtype=torch.float,
                                             device=self.device)
        # dict in the reference's key order (reward_scales order); values are rows of one [R,N] buffer
        self.episode_sums = {name: self._episode_sums_buf[self._sum_row[name]] for name in self.reward_scales.keys()}
        self._extras_buf = torch.zeros(self._num_sum_rows + 4, dtype=torch.float, device=self.device)
        self.episode_stats = torch.zeros(2 * (self._num_sum_rows + 4), dtype=torch.float, device=self.device)
        if self._external_terms:
            self._ext_rewards = torch.zeros(len(self._external_terms), self.num_envs, dtype=torch.float,
                                            device=self.device)
        else:
            self._ext_rewards = None

    def _init_observation_override(self):
        """A subclass may override compute_observations() to append its own columns (the usual way tasks
        derived from the reference add observations: call super(), then torch.cat).  The fused kernel then
        writes the base observation -- 12 + 3*num_dof (+ height scan) columns, noise and clip included --
        into a buffer of its own, and the override runs right after the launch, where the reference calls
        it too (after reset_idx, :133), with self.obs_buf bound to that base observation."""
        self._obs_override = type(self).compute_observations is not LeggedRobot.compute_observations
        self._termination_override = type(self).check_termination is not LeggedRobot.check_termination
        self._callback_override = (type(self)._post_physics_step_callback
                                   is not LeggedRobot._post_physics_step_callback)
        H = self.num_height_points if self.cfg.terrain.measure_heights else 0
        self._kernel_num_obs = 12 + 3 * self.num_dof + H
        self._kernel_obs = None
        if self._obs_override:
            self._kernel_obs = torch.zeros(self.num_envs, self._kernel_num_obs, dtype=torch.float, device=self.device)
        elif self._kernel_num_obs != self.num_obs:
            raise ValueError(
                f"cfg.env.num_observations={self.num_obs} but the observation of this config has "
                f"{self._kernel_num_obs} columns (12 + 3*num_dof + height points); override compute_observations() "
                f"to add columns")

    def _post_physics_step_callback(self):
        """Command resampling, the heading command, the height scan and the pushes (:337-352) happen on the
        device; a subclass override's super()._post_physics_step_callback() finds commands and measured_heights
        already updated for this step (lgk_pre_reward_state) and may go on from there."""

    def check_termination(self):
        """reset_buf = any termination body in contact | time_out (:143-148) is formed on the device; when a
        subclass overrides this method, its super().check_termination() finds both buffers already filled
        (lgk_pre_reward_state) and may OR further conditions into self.reset_buf in place."""

    def compute_observations(self):
        """The observation (:214-231) -- body-frame velocities, gravity, commands, joint state, actions, the
        height scan, noise -- is written by the fused step kernel; by the time a subclass override calls
        super().compute_observations(), self.obs_buf already holds it.  Nothing to do here."""

    def _run_observation_override(self):
        self.obs_buf = self._kernel_obs
        self.compute_observations()
        if self.obs_buf.shape != (self.num_envs, self.num_obs):
            raise ValueError(f"compute_observations() left obs_buf with shape {tuple(self.obs_buf.shape)}, "
                             f"cfg.env.num_observations says {self.num_obs}")
        clip_obs = self.cfg.normalization.clip_observations          # (:103-104)
        self.obs_buf = torch.clip(self.obs_buf, -clip_obs, clip_obs)

    def _check_hooks(self):
        for name in _FUSED_HOOKS:
            if getattr(type(self), name, None) is not getattr(LeggedRobot, name, None):
                raise NotImplementedError(
                    f"{type(self).__name__} overrides {name}(), whose body runs inside the fused step kernel; "
                    f"overriding it is not supported")
        # reset_idx() itself may be overridden (Anymal does, anymal.py:56-60), but step() never calls it: the
        # per-step resets run inside the fused kernel.  The override's extra part must therefore also exist as
        # _after_fused_resets(), which step() calls with the device-side id list in self.reset_ids /
        # self.step_counts[0]; an override without it would silently miss every in-episode reset.
        if (getattr(type(self), "reset_idx") is not LeggedRobot.reset_idx and
                getattr(type(self), "_after_fused_resets") is LeggedRobot._after_fused_resets):
            raise NotImplementedError(
                f"{type(self).__name__} overrides reset_idx() without _after_fused_resets(): step() resets envs "
                f"inside the fused kernel and never calls reset_idx(); put the override's extra per-env clean-up "
                f"into _after_fused_resets() as well (ids: self.reset_ids[:self.step_counts[0]], on the device)")

    # ------------------------------------------------------------------ terrain / envs
    def _create_ground_plane(self):
        p = gymapi.PlaneParams()
        p.normal = gymapi.Vec3(0.0, 0.0, 1.0)
        p.static_friction = self.cfg.terrain.static_friction
        p.dynamic_friction = self.cfg.terrain.dynamic_friction
        p.restitution = self.cfg.terrain.restitution
        self.gym.add_ground(self.sim, p)

    def _terrain_common(self, p):
        p.transform.p.x = -self.terrain.cfg.border_size
        p.transform.p.y = -self.terrain.cfg.border_size
        p.transform.p.z = 0.0
        p.static_friction = self.cfg.terrain.static_friction
        p.dynamic_friction = self.cfg.terrain.dynamic_friction
        p.restitution = self.cfg.terrain.restitution

    def _bind_height_samples(self):
        if getattr(self.terrain, "height_field_dev", None) is not None:      # generated on the device: no upload
            self.height_samples = self.terrain.height_field_dev.view(self.terrain.tot_rows, self.terrain.tot_cols)
            return
        self.height_samples = torch.tensor(self.terrain.heightsamples).view(
            self.terrain.tot_rows, self.terrain.tot_cols).to(self.device)

    def _create_heightfield(self):
        p = gymapi.HeightFieldParams()
        p.column_scale = self.terrain.cfg.horizontal_scale
        p.row_scale = self.terrain.cfg.horizontal_scale
        p.vertical_scale = self.terrain.cfg.vertical_scale
        p.nbRows = self.terrain.tot_cols
        p.nbColumns = self.terrain.tot_rows
        self._terrain_common(p)
        self.gym.add_heightfield(self.sim, self.terrain.heightsamples, p)
        self._bind_height_samples()

    def _create_trimesh(self):
        p = gymapi.TriangleMeshParams()
        p.nb_vertices = self.terrain.vertices.shape[0]
        p.nb_triangles = self.terrain.triangles.shape[0]
        self._terrain_common(p)
        self.gym.add_triangle_mesh(self.sim, self.terrain.vertices.flatten(order='C'),
                                   self.terrain.triangles.flatten(order='C'), p)
        self._bind_height_samples()

    def _create_envs(self):
        """Load the asset, create one actor per env, resolve body index lists (:645-740)."""
        asset_path = self.cfg.asset.file.format(LEGGED_GYM_ROOT_DIR=LEGGED_GYM_ROOT_DIR)
        asset_root, asset_file = os.path.dirname(asset_path), os.path.basename(asset_path)
        opt = gymapi.AssetOptions()
        for k in ("default_dof_drive_mode", "collapse_fixed_joints", "replace_cylinder_with_capsule",
                  "flip_visual_attachments", "fix_base_link", "density", "angular_damping", "linear_damping",
                  "max_angular_velocity", "max_linear_velocity", "armature", "thickness", "disable_gravity"):
            setattr(opt, k, getattr(self.cfg.asset, k))
        robot_asset = self.gym.load_asset(self.sim, asset_root, asset_file, opt)
        self.num_dof = self.gym.get_asset_dof_count(robot_asset)
        dof_props_asset = self.gym.get_asset_dof_properties(robot_asset)
        rigid_shape_props_asset = self.gym.get_asset_rigid_shape_properties(robot_asset)
        body_names = self.gym.get_asset_rigid_body_names(robot_asset)
        self.dof_names = self.gym.get_asset_dof_names(robot_asset)
        self.num_bodies = len(body_names)
        self.num_dofs = len(self.dof_names)
        feet_names = [s for s in body_names if self.cfg.asset.foot_name in s]
        penalized = [s for n in self.cfg.asset.penalize_contacts_on for s in body_names if n in s]
        terminating = [s for n in self.cfg.asset.terminate_after_contacts_on for s in body_names if n in s]

        for i in range(self.num_dof):
            for dof_name in self.cfg.control.stiffness.keys():
                if dof_name in self.dof_names[i]:
                    dof_props_asset['driveMode'][i] = gymapi.DOF_MODE_POS
                    dof_props_asset['stiffness'][i] = self.cfg.control.stiffness[dof_name]
                    dof_props_asset['damping'][i] = self.cfg.control.damping[dof_name]

        ist = self.cfg.init_state
        self.base_init_state = torch.tensor(ist.pos + ist.rot + ist.lin_vel + ist.ang_vel, dtype=torch.float,
                                            device=self.device, requires_grad=False)
        self._get_env_origins()
        lower, upper = gymapi.Vec3(0., 0., 0.), gymapi.Vec3(0., 0., 0.)
        start = gymapi.Transform()
        origins_host = self.env_origins.cpu()
        jitter = (2.0 * torch.rand(self.num_envs, 2) - 1.0)
        per_row = int(np.sqrt(self.num_envs))
        self.actor_handles, self.envs = [], []
        for i in range(self.num_envs):
            env_handle = self.gym.create_env(self.sim, lower, upper, per_row)
            pos = origins_host[i].clone()
            pos[:2] += jitter[i]
            start.p = gymapi.Vec3(*pos.tolist())
            shape_props = self._process_rigid_shape_props(rigid_shape_props_asset, i)
            self.gym.set_asset_rigid_shape_properties(robot_asset, shape_props)
            actor = self.gym.create_actor(env_handle, robot_asset, start, self.cfg.asset.name, i,
                                          self.cfg.asset.self_collisions, 0)
            dof_props = self._process_dof_props(dof_props_asset, i)
            self.gym.set_actor_dof_properties(env_handle, actor, dof_props)
            self.gym.enable_actor_dof_force_sensors(env_handle, actor)
            body_props = self.gym.get_actor_rigid_body_properties(env_handle, actor)
            body_props = self._process_rigid_body_props(body_props, i)
            self.gym.set_actor_rigid_body_properties(env_handle, actor, body_props, recomputeInertia=True)
            self.envs.append(env_handle)
            self.actor_handles.append(actor)

        def handles(names):
            out = torch.zeros(len(names), dtype=torch.long, device=self.device, requires_grad=False)
            for k, n in enumerate(names):
                out[k] = self.gym.find_actor_rigid_body_handle(self.envs[0], self.actor_handles[0], n)
            return out
        self.feet_indices = handles(feet_names)
        self.penalised_contact_indices = handles(penalized)
        self.termination_contact_indices = handles(terminating)

    def _get_env_origins(self):
        """Terrain platforms on rough terrain, a grid otherwise (:742-767).  With env sharding
        (cfg.env.shard_rank / shard_world) terrain types follow the GLOBAL env index."""
        N = self.num_envs
        rank = int(getattr(self.cfg.env, "shard_rank", 0))
        world = int(getattr(self.cfg.env, "shard_world", 1))
        self.env_id_offset = rank * N
        n_total = N * world
        if self.cfg.terrain.mesh_type in ["heightfield", "trimesh"]:
            self.custom_origins = True
            self.env_origins = torch.zeros(N, 3, device=self.device, requires_grad=False)
            max_init_level = self.cfg.terrain.max_init_terrain_level
            if not self.cfg.terrain.curriculum:
                max_init_level = self.cfg.terrain.num_rows - 1
            self.terrain_levels = torch.randint(0, max_init_level + 1, (N,), device=self.device)
            gid = torch.arange(N, device=self.device) + self.env_id_offset
            self.terrain_types = torch.div(gid, (n_total / self.cfg.terrain.num_cols), rounding_mode='floor').to(torch.long)
            self.max_terrain_level = self.cfg.terrain.num_rows
            self.terrain_origins = torch.from_numpy(self.terrain.env_origins).to(self.device).to(torch.float)
            self.env_origins[:] = self.terrain_origins[self.terrain_levels, self.terrain_types]
        else:
            self.custom_origins = False
            self.env_origins = torch.zeros(N, 3, device=self.device, requires_grad=False)
            num_cols = np.floor(np.sqrt(N))
            num_rows = np.ceil(N / num_cols)
            xx, yy = torch.meshgrid(torch.arange(num_rows), torch.arange(num_cols), indexing="ij")
            spacing = self.cfg.env.env_spacing
            self.env_origins[:, 0] = spacing * xx.flatten()[:N]
            self.env_origins[:, 1] = spacing * yy.flatten()[:N]
            self.env_origins[:, 2] = 0.

    def _parse_cfg(self, cfg):
        self.dt = self.cfg.control.decimation * self.sim_params.dt
        self.obs_scales = self.cfg.normalization.obs_scales
        self.reward_scales = class_to_dict(self.cfg.rewards.scales)
        self.command_ranges = class_to_dict(self.cfg.commands.ranges)
        if self.cfg.terrain.mesh_type not in ['heightfield', 'trimesh']:
            self.cfg.terrain.curriculum = False
        self.max_episode_length_s = self.cfg.env.episode_length_s
        self.max_episode_length = np.ceil(self.max_episode_length_s / self.dt)
        self.cfg.domain_rand.push_interval = np.ceil(self.cfg.domain_rand.push_interval_s / self.dt)

    def _init_height_points(self):
        """Scan offsets in the base frame, x-major (:802-816). Stored once and exposed as an
        expanded [N,H,3] view (the reference materialises N copies)."""
        y = torch.tensor(self.cfg.terrain.measured_points_y, device=self.device, requires_grad=False)
        x = torch.tensor(self.cfg.terrain.measured_points_x, device=self.device, requires_grad=False)
        grid_x, grid_y = torch.meshgrid(x, y, indexing="ij")
        self.num_height_points = grid_x.numel()
        pts = torch.zeros(self.num_height_points, 3, device=self.device, requires_grad=False)
        pts[:, 0] = grid_x.flatten()
        pts[:, 1] = grid_y.flatten()
        self._height_points_xy = pts[:, :2].contiguous()
        return pts.unsqueeze(0).expand(self.num_envs, -1, -1)

    def _change_cmds(self, vx, vy, vang):
        self.commands[:, 0] = vx
        self.commands[:, 1] = vy
        self.commands[:, 2] = vang

    # ------------------------------------------------------------------ kernel tables
    def _kernels(self):
        if self._lib is None:
            if not str(self.device).startswith("cuda"):
                raise RuntimeError("LeggedRobot.step() needs a CUDA device (sm_100a kernels); there is no CPU path")
            self._lib = _lgk.load()
            self.refresh_bindings()
        return self._lib

    def refresh_bindings(self):
        """(Re)build the constant block and the device-pointer table handed to the kernels.  Call it
        after re-binding (not after writing into) any public tensor."""
        self._consts = self.build_consts()
        self._build_terrain_tables()
        self._bufs = self.build_buffers()
        self._ranges_cache = None
        self._ranges_split = False
        self._sync_host_scalars()
        self._ensure_dev_counter(self.common_step_counter)

    def build_consts(self, rng_mode=_lgk.RNG_PHILOX, index_math=None) -> "_lgk.LgkConsts":
        cfg = self.cfg
        c = _lgk.LgkConsts()
        c.struct_size = C.sizeof(_lgk.LgkConsts)
        c.abi_version = _lgk.LGK_ABI_VERSION
        D = self.num_dof
        if D > _lgk.LGK_MAX_DOF:
            raise ValueError(f"num_dof={D} exceeds LGK_MAX_DOF={_lgk.LGK_MAX_DOF}")
        H = self.num_height_points if cfg.terrain.measure_heights else 0
        c.num_envs, c.num_dof, c.num_bodies = self.num_envs, D, self.num_bodies
        c.num_feet, c.num_pen, c.num_term = len(self.feet_indices), len(self.penalised_contact_indices), \
            len(self.termination_contact_indices)
        c.num_height_points, c.num_obs = H, self._kernel_num_obs
        c.num_rewards = len(self.reward_names)
        c.num_commands = cfg.commands.num_commands
        c.env_id_offset = getattr(self, "env_id_offset", 0)
        flags = 0
        flags |= _lgk.F_HEADING_COMMAND if cfg.commands.heading_command else 0
        flags |= _lgk.F_MEASURE_HEIGHTS if cfg.terrain.measure_heights else 0
        flags |= _lgk.F_TERRAIN_PLANE if cfg.terrain.mesh_type == 'plane' else 0
        flags |= _lgk.F_ADD_NOISE if self.add_noise else 0
        flags |= _lgk.F_PUSH_ROBOTS if cfg.domain_rand.push_robots else 0
        flags |= _lgk.F_TERRAIN_CURRICULUM if cfg.terrain.curriculum else 0
        flags |= _lgk.F_CUSTOM_ORIGINS if self.custom_origins else 0
        flags |= _lgk.F_ONLY_POSITIVE if cfg.rewards.only_positive_rewards else 0
        flags |= _lgk.F_HAS_TERMINATION if "termination" in self.reward_scales else 0
        flags |= _lgk.F_INIT_DONE if self.init_done else 0
        flags |= _lgk.F_FILTERED_CONTACTS if getattr(cfg.rewards, "filtered_feet_contact", False) else 0
        c.flags = flags
        if cfg.terrain.measure_heights and cfg.terrain.mesh_type == 'none':
            raise NameError("Can't measure height with terrain mesh type 'none'")   # (:833-834)
        if cfg.control.control_type not in _lgk.CTRL:
            raise NameError(f"Unknown controller type: {cfg.control.control_type}")   # (:391)
        c.control_type = _lgk.CTRL[cfg.control.control_type]
        if index_math is None:
            index_math = int(getattr(cfg.env, "index_math", _lgk.INDEX_MATH_TORCH_CUDA))
        c.index_math = index_math
        c.rng_mode = rng_mode
        c.envs_per_cta = int(getattr(cfg.env, "envs_per_cta", 0))
        c.dt, c.sim_dt = self.dt, self.sim_params.dt
        c.max_episode_length = int(self.max_episode_length)
        c.resample_interval = int(cfg.commands.resampling_time / self.dt)
        c.push_interval = int(cfg.domain_rand.push_interval)
        c.max_episode_length_s = self.max_episode_length_s
        c.max_push_vel_xy = cfg.domain_rand.max_push_vel_xy
        c.clip_actions = cfg.normalization.clip_actions
        c.clip_obs = cfg.normalization.clip_observations
        s = self.obs_scales
        c.obs_scale_lin_vel, c.obs_scale_ang_vel = s.lin_vel, s.ang_vel
        c.obs_scale_dof_pos, c.obs_scale_dof_vel, c.obs_scale_height = s.dof_pos, s.dof_vel, s.height_measurements
        for i, v in enumerate(self.commands_scale.tolist()):
            c.commands_scale[i] = v
        c.action_scale = cfg.control.action_scale

        def fill(dst, t):
            for i, v in enumerate(t.detach().flatten().cpu().tolist()):
                dst[i] = v
        fill(c.p_gains, self.p_gains)
        fill(c.d_gains, self.d_gains)
        fill(c.torque_limits, self.torque_limits)
        fill(c.default_dof_pos, self.default_dof_pos)
        fill(c.dof_pos_lo, self.dof_pos_limits[:, 0])
        fill(c.dof_pos_hi, self.dof_pos_limits[:, 1])
        fill(c.dof_vel_limits, self.dof_vel_limits)
        r = cfg.rewards
        c.soft_dof_vel_limit, c.soft_torque_limit = r.soft_dof_vel_limit, r.soft_torque_limit
        c.base_height_target, c.tracking_sigma, c.max_contact_force = r.base_height_target, r.tracking_sigma, \
            r.max_contact_force
        c.termination_scale = self.reward_scales.get("termination", 0.0)
        if len(self.reward_names) + (1 if "termination" in self.reward_scales else 0) > _lgk.LGK_MAX_REWARDS:
            raise ValueError("too many active reward terms")
        for i, (name, tid) in enumerate(zip(self.reward_names, self._reward_term_ids)):
            c.rewards[i].term = tid
            c.rewards[i].scale = self.reward_scales[name]
        fill(c.feet_idx, self.feet_indices)
        fill(c.pen_idx, self.penalised_contact_indices)
        fill(c.term_idx, self.termination_contact_indices)
        hips = [i for i in (0, 3, 6, 9) if i < D]     # hard-coded quadruped hips (:966)
        for i, v in enumerate(hips):
            c.hip_idx[i] = v
        c.num_hip = len(hips)
        fill(c.base_init_state, self.base_init_state)
        t = cfg.terrain
        c.border_size, c.horizontal_scale, c.vertical_scale = t.border_size, t.horizontal_scale, t.vertical_scale
        if self.height_samples is not None:
            c.half_env_length = self.terrain.env_length / 2
            c.terrain_rows, c.terrain_cols = self.height_samples.shape
            c.terrain_pitch = (self.height_samples.shape[1] + 7) // 8 * 8
            c.max_terrain_level = self.max_terrain_level
            c.num_terrain_types = cfg.terrain.num_cols
        seed = int(getattr(cfg, "seed", 1))
        c.seed = ((seed * 0x9E3779B97F4A7C15) ^ (c.env_id_offset * 0x632BE59BD9B4E019)) & ((1 << 64) - 1)
        return c

    def _build_terrain_tables(self):
        """height_min3[x,y] = min(h[x,y], h[x+1,y], h[x,y+1]): the three gathers + two mins of
        _get_heights (:848-852) folded into a table once, since the terrain never changes."""
        self._height_min3 = None
        if self.height_samples is None or not str(self.device).startswith("cuda"):
            return
        rows, cols = self.height_samples.shape
        pitch = self._consts.terrain_pitch
        hs = self.height_samples.contiguous()
        self._height_min3 = torch.empty(rows, pitch, dtype=torch.int16, device=self.device)
        _lgk.check("lgk_build_height_min3", self._lib.lgk_build_height_min3(
            _lgk.ptr(hs), rows, cols, _lgk.ptr(self._height_min3), pitch, _lgk.current_stream(self.device)))

    def build_buffers(self, taps: Optional[dict] = None) -> "_lgk.LgkBuffers":
        b = _lgk.LgkBuffers()
        P = _lgk.ptr
        N = self.num_envs
        if not hasattr(self, "_workspace"):
            nbytes = self._lib.lgk_workspace_bytes(N, self._num_sum_rows)
            self._workspace = torch.zeros((nbytes + 15) // 16 * 2, dtype=torch.int64, device=self.device)
            self._command_ranges_dev = torch.zeros(16, dtype=torch.float, device=self.device)
        for name in ("root_states", "dof_state", "contact_forces", "torques", "actions", "last_actions",
                     "last_dof_vel", "last_root_vel", "commands", "feet_air_time", "env_origins", "obs_buf",
                     "rew_buf", "reset_buf", "time_out_buf", "base_lin_vel", "base_ang_vel", "projected_gravity",
                     "reset_ids", "resample_ids", "noise_scale_vec"):
            t = getattr(self, name)
            if not t.is_contiguous():
                raise ValueError(f"{name} must be contiguous")
            setattr(b, name, P(t))
        if self._obs_override:
            b.obs_buf = P(self._kernel_obs)
        b.episode_sums = P(self._episode_sums_buf)
        b.episode_length_buf = P(self._episode_length_buf)
        b.terrain_levels = P(getattr(self, "terrain_levels", None))
        b.terrain_types = P(getattr(self, "terrain_types", None))
        b.measured_heights = P(self.measured_heights) if torch.is_tensor(self.measured_heights) else None
        b.counts = P(self.step_counts)
        b.episode_stats = P(self.episode_stats)
        b.extras = P(self._extras_buf)
        b.ext_rewards = P(self._ext_rewards)
        b.height_min3 = P(getattr(self, "_height_min3", None))
        b.height_points_xy = P(getattr(self, "_height_points_xy", None))
        b.terrain_origins = P(getattr(self, "terrain_origins", None))
        b.command_ranges = P(self._command_ranges_dev)
        b.last_contacts = P(self.last_contacts)
        b.workspace = P(self._workspace)
        if taps:
            for k, t in taps.items():
                setattr(b, "tap_" + k, P(t))
        return b

    def use_rng_taps(self, taps: Optional[dict], index_math: Optional[int] = None):
        """Parity mode: draw every random number of the step from caller-provided per-env uniform
        tensors (keys cmd_callback, cmd_reset, push, dof, root_xy, root_vel, noise, level) instead of
        the in-kernel Philox generator; `None` switches back."""
        self._kernels()
        if taps is None:
            self._taps = None
            self._consts.rng_mode = _lgk.RNG_PHILOX
            self._bufs = self.build_buffers()
        else:
            self._taps = {k: v.to(self.device).contiguous() for k, v in taps.items()}
            self._taps["level"] = self._taps["level"].to(torch.long)
            self._consts.rng_mode = _lgk.RNG_TAPS
            self._bufs = self.build_buffers(self._taps)
        if index_math is not None:
            self._consts.index_math = index_math

    def bind_sim_tensors(self, root_states=None, dof_state=None, contact_forces=None, torques=None):
        """Point the env at other storage for the simulator-owned tensors (a backend that hands out a
        different buffer per frame, or pre-generated frames resident in HBM) without copying."""
        self._kernels()
        N, D = self.num_envs, self.num_dof
        if root_states is not None:
            self.root_states = root_states.view(N, 13)
            self.base_quat = self.root_states[:, 3:7]
        if dof_state is not None:
            self.dof_state = dof_state.view(N * D, 2)
            self.dof_pos = self.dof_state.view(N, D, 2)[..., 0]
            self.dof_vel = self.dof_state.view(N, D, 2)[..., 1]
        if contact_forces is not None:
            self.contact_forces = contact_forces.view(N, -1, 3)
        if torques is not None:
            self.torques = torques.view(N, self.num_actions)
        P = _lgk.ptr
        self._bufs.root_states, self._bufs.dof_state = P(self.root_states), P(self.dof_state)
        self._bufs.contact_forces, self._bufs.torques = P(self.contact_forces), P(self.torques)

    def bind_output_tensors(self, obs_buf=None, rew_buf=None, reset_buf=None):
        """Let the fused kernel write observations / rewards / resets straight into the caller's storage
        (e.g. slot t of a [T,N,O] rollout buffer) instead of into env-owned buffers that the caller then
        copies (the reference's runner does `storage.observations[t].copy_(obs)`,
        a second pass over the largest tensor of the step).  Each tensor must be contiguous with the
        same element count and width as the buffer it replaces; `reset_buf` may be bool or uint8.
        step() returns the bound tensors."""
        self._kernels()
        N = self.num_envs
        if obs_buf is not None:
            if self._obs_override:
                raise ValueError("obs_buf is produced by this task's compute_observations() override, not by the kernel")
            if obs_buf.dtype != torch.float or obs_buf.numel() != N * self.num_obs or not obs_buf.is_contiguous():
                raise ValueError(f"obs_buf must be contiguous float32 with {N}x{self.num_obs} elements")
            self.obs_buf = obs_buf.view(N, self.num_obs)
            self._bufs.obs_buf = _lgk.ptr(self.obs_buf)
        if rew_buf is not None:
            if rew_buf.dtype != torch.float or rew_buf.numel() != N or not rew_buf.is_contiguous():
                raise ValueError(f"rew_buf must be contiguous float32 with {N} elements")
            self.rew_buf = rew_buf.view(N)
            self._bufs.rew_buf = _lgk.ptr(self.rew_buf)
        if reset_buf is not None:
            if reset_buf.dtype not in (torch.bool, torch.uint8) or reset_buf.numel() != N or not reset_buf.is_contiguous():
                raise ValueError(f"reset_buf must be contiguous bool / uint8 with {N} elements")
            self.reset_buf = reset_buf.view(N)
            self._bufs.reset_buf = _lgk.ptr(self.reset_buf)

    # ------------------------------------------------------------------ env-state checkpoint
    # the reference resumes the policy only (utils/helpers.py:103-125,
    # task_registry.py:162-167) and restarts every env from scratch.  Here the whole env state is a fixed
    # set of device buffers plus three host scalars, and the in-kernel RNG is counter-based -- keyed by
    # (seed, common_step_counter, env id, site) -- so restoring the counter restores the random stream.
    _STATE_TENSORS = ("root_states", "dof_state", "actions", "last_actions", "last_dof_vel", "last_root_vel",
                      "commands", "feet_air_time", "env_origins", "obs_buf", "rew_buf", "reset_buf", "time_out_buf",
                      "base_lin_vel", "base_ang_vel", "projected_gravity", "_episode_sums_buf", "episode_stats",
                      "_extras_buf", "_episode_length_buf", "terrain_levels", "last_contacts")

    def _extra_state_tensors(self) -> Dict[str, torch.Tensor]:
        """Per-env state a subclass adds (actuator histories, LSTM state)."""
        return {}

    def state_dict(self) -> dict:
        """Everything needed to continue this env bit-for-bit: host copies of the state buffers, the step
        counter, the (curriculum-mutable) command ranges and the seed."""
        tensors = {n: getattr(self, n).detach().cpu().clone() for n in self._STATE_TENSORS
                   if torch.is_tensor(getattr(self, n, None))}
        tensors.update({"extra." + k: v.detach().cpu().clone() for k, v in self._extra_state_tensors().items()})
        return {"version": 1, "task": type(self).__name__, "num_envs": self.num_envs, "num_obs": self.num_obs,
                "seed": int(getattr(self.cfg, "seed", 1)), "env_id_offset": int(getattr(self, "env_id_offset", 0)),
                "common_step_counter": int(self.common_step_counter),
                "command_ranges": {k: [float(x) for x in v] for k, v in self.command_ranges.items()},
                "reward_names": list(self.reward_names), "tensors": tensors}

    def load_state_dict(self, sd: dict):
        """Restore a state_dict() into this env's existing buffers (addresses stay as the kernels know
        them).  The env must have been built from the same config."""
        if sd.get("version") != 1:
            raise ValueError(f"unknown env checkpoint version {sd.get('version')}")
        for key in ("num_envs", "num_obs"):
            if sd[key] != getattr(self, key):
                raise ValueError(f"checkpoint {key}={sd[key]} != env {key}={getattr(self, key)}")
        if list(sd["reward_names"]) != list(self.reward_names):
            raise ValueError("checkpoint was taken with a different reward set")
        extra = self._extra_state_tensors()
        for name, t in sd["tensors"].items():
            dst = extra.get(name[6:]) if name.startswith("extra.") else getattr(self, name, None)
            if not torch.is_tensor(dst) or dst.numel() != t.numel() or dst.dtype != t.dtype:
                raise ValueError(f"checkpoint tensor {name} does not fit this env")
            dst.copy_(t.view(dst.shape))
        for k, v in sd["command_ranges"].items():
            self.command_ranges[k][0], self.command_ranges[k][1] = v[0], v[1]
        self.common_step_counter = int(sd["common_step_counter"])
        self._dev_step_counter = -1          # re-seed the device copy on the next step
        self._ranges_cache = None
        if self._lib is not None:
            self._sync_host_scalars()

    def _ensure_dev_counter(self, want: int):
        """Device copy of common_step_counter: the step kernel increments it itself, so it is
        re-seeded only when the host value was changed from outside."""
        if self._dev_step_counter != want:
            _lgk.check("lgk_set_step_counter", self._lib.lgk_set_step_counter(
                _lgk.ptr(self._workspace), int(want), _lgk.current_stream(self.device)))
            self._dev_step_counter = want

    def _sync_host_scalars(self, reset_half_only: bool = False):
        """Mirror host-side scalars the kernels read from device memory: command_ranges (a mutable
        dict in the reference, :473-474) and common_step_counter.  The device copy holds the ranges twice:
        for the callback's resample and for reset_idx's; `reset_half_only` updates the latter alone (the
        command curriculum changes the dict between the two on its gated step)."""
        cr = self.command_ranges
        key = tuple(float(v) for k in ("lin_vel_x", "lin_vel_y", "ang_vel_yaw", "heading") for v in cr[k])
        if key != self._ranges_cache or (self._ranges_split and not reset_half_only):
            vals = []
            for lo, hi in zip(key[0::2], key[1::2]):
                vals += [lo, hi - lo]         # (lower, upper - lower) as torch_rand_float forms them
            t = torch.tensor(vals, dtype=torch.float)
            self._command_ranges_dev[8:].copy_(t)
            if not reset_half_only:
                self._command_ranges_dev[:8].copy_(t)
            self._ranges_split = reset_half_only
            self._ranges_cache = key
        flag = _lgk.F_INIT_DONE if self.init_done else 0
        self._consts.flags = (self._consts.flags & ~_lgk.F_INIT_DONE) | flag
