"""Rollout storage the step kernel writes into directly.

The reference hands `obs, rewards, dones` of every `env.step()` to rsl_rl, whose RolloutStorage copies
them into `[T, N, ...]` tensors (`task_registry.py:160` -> OnPolicyRunner; `storage.observations[t].copy_`):
one more read + write of the largest tensor of the step (940 B per env at 235 observations, about half
of the step's algorithmic bytes).  Here the storage owns the memory and the env's output pointers are
moved to slot t before each step, so the fused kernel's stores are the only pass.

Layout follows rsl_rl's RolloutStorage: `observations [T(+1), N, O]`, `rewards [T, N, 1]`,
`dones [T, N, 1]` (bool).  observations[t] is the observation the policy acted on at transition t, i.e.
the output of step t-1; slot T receives the output of the last step and seeds the next rollout.
"""
from __future__ import annotations

import torch


class RolloutStorageWriter:
    def __init__(self, env, num_transitions_per_env: int):
        self.env = env
        self.T = int(num_transitions_per_env)
        N, O, dev = env.num_envs, env.num_obs, env.device
        self.observations = torch.zeros(self.T + 1, N, O, dtype=torch.float, device=dev)
        self.rewards = torch.zeros(self.T, N, 1, dtype=torch.float, device=dev)
        self.dones = torch.zeros(self.T, N, 1, dtype=torch.bool, device=dev)
        self.step_idx = 0
        self.observations[0].copy_(env.get_observations())

    def step(self, actions):
        """env.step(actions) with outputs landing in observations[t+1], rewards[t], dones[t]."""
        if self.step_idx >= self.T:
            raise AssertionError("Rollout buffer overflow")     # same failure as rsl_rl's add_transitions
        t = self.step_idx
        N = self.env.num_envs
        self.env.bind_output_tensors(obs_buf=self.observations[t + 1], rew_buf=self.rewards[t].view(N),
                                     reset_buf=self.dones[t].view(N))
        out = self.env.step(actions)
        self.step_idx += 1
        return out

    def clear(self):
        """Start the next rollout from the last observation (one [N,O] copy per T steps)."""
        self.observations[0].copy_(self.observations[self.T])
        self.step_idx = 0
