"""Buffer contract of a vectorised task (reference ``envs/base/base_task.py:38-144``).

Allocates the rsl_rl ``VecEnv`` surface -- ``obs_buf``, ``rew_buf``, ``reset_buf``,
``episode_length_buf``, ``time_out_buf``, ``privileged_obs_buf``, ``extras`` -- creates the
simulator and offers ``reset()`` / ``get_observations()``.  The viewer / ``render()`` half of the
reference class is left out: there is no graphics stack on the target machines.

Simulator binding: the real Isaac Gym is used when it is importable, otherwise the headless
backend in :mod:`legged_gym_b200.sim` (same tensor API, no physics).
"""
from __future__ import annotations

import torch

from legged_gym_b200.sim.backend import REAL_ISAACGYM as HAVE_ISAACGYM  # real simulator if present, else the stand-in
from legged_gym_b200.sim.backend import gymapi, gymutil


class BaseTask:
    def __init__(self, cfg, sim_params, physics_engine, sim_device, headless):
        self.gym = gymapi.acquire_gym()
        self.sim_params = sim_params
        self.physics_engine = physics_engine
        self.sim_device = sim_device
        sim_device_type, self.sim_device_id = gymutil.parse_device_str(self.sim_device)
        self.headless = headless

        # the env lives on the GPU only when the simulator runs its tensor pipeline there
        if sim_device_type == 'cuda' and sim_params.use_gpu_pipeline:
            self.device = self.sim_device
        else:
            self.device = 'cpu'
        self.graphics_device_id = -1 if self.headless else self.sim_device_id

        self.num_envs = cfg.env.num_envs
        self.num_obs = cfg.env.num_observations
        self.num_privileged_obs = cfg.env.num_privileged_obs
        self.num_actions = cfg.env.num_actions

        f32 = dict(device=self.device, dtype=torch.float)
        self.obs_buf = torch.zeros(self.num_envs, self.num_obs, **f32)
        self.rew_buf = torch.zeros(self.num_envs, **f32)
        # reference allocates int64 ones and re-binds a bool tensor on the first
        # check_termination (base_task.py:72, legged_robot.py:146); here it is bool throughout
        self.reset_buf = torch.ones(self.num_envs, device=self.device, dtype=torch.bool)
        self.episode_length_buf = torch.zeros(self.num_envs, device=self.device, dtype=torch.long)
        self.time_out_buf = torch.zeros(self.num_envs, device=self.device, dtype=torch.bool)
        if self.num_privileged_obs is not None:
            self.privileged_obs_buf = torch.zeros(self.num_envs, self.num_privileged_obs, **f32)
        else:
            self.privileged_obs_buf = None
        self.extras = {}

        self.create_sim()
        self.gym.prepare_sim(self.sim)

        self.enable_viewer_sync = True
        self.viewer = None
        if not self.headless:
            raise RuntimeError("no graphics stack: construct tasks with headless=True")

    def get_observations(self):
        return self.obs_buf

    def get_privileged_observations(self):
        return self.privileged_obs_buf

    def reset_idx(self, env_ids):
        raise NotImplementedError

    def reset(self):
        """Reset all robots, then take one zero-action step (base_task.py:111-115)."""
        self.reset_idx(torch.arange(self.num_envs, device=self.device))
        obs, privileged_obs, _, _, _ = self.step(
            torch.zeros(self.num_envs, self.num_actions, device=self.device, requires_grad=False))
        return obs, privileged_obs

    def step(self, actions):
        raise NotImplementedError

    def render(self, sync_frame_time=True):
        return None   # viewer-only in the reference; nothing to do headless
