"""Seeded synthetic simulator state for tests and the benchmark.

No physics runs anywhere in this repository: the four simulator tensors and the policy actions are
drawn from fixed distributions chosen so that every branch of the step fires at a realistic rate
(~1 % contact terminations, soft joint limits crossed by ~5 % of joints, feet in contact half of the
time, ...).  Works on any env object exposing the reference's attribute names.
"""
from __future__ import annotations

from typing import Dict, List

import torch


def make_frames(env, num_frames: int, seed: int, device=None) -> List[Dict[str, torch.Tensor]]:
    """Returns `num_frames` dicts with root_states, dof_state, contact_forces, torques, actions (CPU
    generation, then moved to `device`)."""
    g = torch.Generator().manual_seed(seed)
    N, D, B = env.num_envs, env.num_dof, env.num_bodies
    device = device or env.device
    origins = env.env_origins.detach().cpu().float()
    default = env.default_dof_pos.detach().cpu().float().view(1, D)
    feet = env.feet_indices.cpu().tolist()
    pen = env.penalised_contact_indices.cpu().tolist()
    term = env.termination_contact_indices.cpu().tolist()
    hs = env.height_samples.cpu() if getattr(env, "height_samples", None) is not None else None
    tcfg = env.cfg.terrain
    init_z = float(env.cfg.init_state.pos[2])
    eff = env.torque_limits.detach().cpu().float().view(1, D)

    def randn(*s):
        return torch.randn(*s, generator=g)

    def rand(*s):
        return torch.rand(*s, generator=g)

    frames = []
    for _ in range(num_frames):
        root = torch.zeros(N, 13)
        root[:, :2] = origins[:, :2] + (2 * rand(N, 2) - 1)
        if hs is not None:
            ix = ((root[:, 0] + tcfg.border_size) / tcfg.horizontal_scale).long().clamp(0, hs.shape[0] - 1)
            iy = ((root[:, 1] + tcfg.border_size) / tcfg.horizontal_scale).long().clamp(0, hs.shape[1] - 1)
            ground = hs[ix, iy].float() * tcfg.vertical_scale
        else:
            ground = torch.zeros(N)
        root[:, 2] = ground + init_z + 0.02 * randn(N)
        q = torch.cat([0.1 * randn(N, 2), randn(N, 2)], dim=1)     # mostly yaw, small roll / pitch
        root[:, 3:7] = q / q.norm(dim=1, keepdim=True).clamp(min=1e-6)
        root[:, 7:13] = 0.5 * randn(N, 6)
        dof = torch.zeros(N, D, 2)
        dof[..., 0] = default + 0.2 * randn(N, D)
        dof[..., 1] = 1.5 * randn(N, D)
        cf = torch.zeros(N, B, 3)
        if feet:
            on = (rand(N, len(feet)) < 0.5).float()
            cf[:, feet, 2] = on * 60.0 * rand(N, len(feet))
            cf[:, feet, :2] = on.unsqueeze(-1) * 3.0 * randn(N, len(feet), 2)
        for ids, p, lo, hi in ((pen, 0.02, 0.1, 5.0), (term, 0.01, 1.0, 10.0)):
            if ids:
                hit = (rand(N, len(ids)) < p).float()
                d = randn(N, len(ids), 3)
                d = d / d.norm(dim=-1, keepdim=True).clamp(min=1e-6)
                mag = lo + (hi - lo) * rand(N, len(ids))
                cf[:, ids, :] = cf[:, ids, :] + (hit * mag).unsqueeze(-1) * d
        tq = torch.minimum(torch.maximum(5.0 * randn(N, D), -eff), eff)
        actions = randn(N, D).clamp(-1, 1)
        frames.append(dict(root_states=root.to(device), dof_state=dof.view(N * D, 2).to(device),
                           contact_forces=cf.view(N * B, 3).to(device), torques=tq.to(device),
                           actions=actions.to(device)))
    return frames


def make_taps(env, seed: int, device=None) -> Dict[str, torch.Tensor]:
    """One set of per-env uniform taps (reference_loader.TapHarness / LgkBuffers.tap_*)."""
    g = torch.Generator().manual_seed(seed)
    N, D, O = env.num_envs, env.num_dof, env.num_obs
    device = device or env.device
    lv = int(getattr(env, "max_terrain_level", 1) or 1)
    taps = dict(cmd_callback=torch.rand(N, 3, generator=g), cmd_reset=torch.rand(N, 3, generator=g),
                push=torch.rand(N, 2, generator=g), dof=torch.rand(N, D, generator=g),
                root_xy=torch.rand(N, 2, generator=g), root_vel=torch.rand(N, 6, generator=g),
                noise=torch.rand(N, O, generator=g), level=torch.randint(0, lv, (N,), generator=g))
    return {k: v.to(device) for k, v in taps.items()}


def load_frame(env, frame: Dict[str, torch.Tensor]):
    """Write a synthetic frame into the simulator-owned tensors (what `simulate` + `refresh` would do)."""
    env.root_states.copy_(frame["root_states"].view_as(env.root_states))
    env.dof_state.copy_(frame["dof_state"].view_as(env.dof_state))
    env.contact_forces.copy_(frame["contact_forces"].view_as(env.contact_forces))
    env.torques.copy_(frame["torques"].view_as(env.torques))
