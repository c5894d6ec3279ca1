"""Torch helpers with the reference's names (``utils/math.py:38-56``) for user hooks.

The kernels do this arithmetic themselves; these exist so subclass code written against the
reference (``quat_apply_yaw``, ``wrap_to_pi``) keeps importing.
"""
import numpy as np
import torch


def normalize(x, eps: float = 1e-9):
    return x / x.norm(p=2, dim=-1).clamp(min=eps).unsqueeze(-1)


def quat_apply(q, v):
    shape = v.shape
    q = q.reshape(-1, 4)
    v = v.reshape(-1, 3)
    u = q[:, :3]
    t = 2 * torch.cross(u, v, dim=-1)
    return (v + q[:, 3:] * t + torch.cross(u, t, dim=-1)).view(shape)


def quat_rotate_inverse(q, v):
    w = q[:, 3:4]
    u = q[:, :3]
    return v * (2.0 * w * w - 1.0) - 2.0 * w * torch.cross(u, v, dim=-1) + 2.0 * u * (u * v).sum(-1, keepdim=True)


def quat_apply_yaw(quat, vec):
    q = quat.clone().view(-1, 4)
    q[:, :2] = 0.
    return quat_apply(normalize(q), vec)


def wrap_to_pi(angles):
    angles %= 2 * np.pi
    angles -= 2 * np.pi * (angles > np.pi)
    return angles


def torch_rand_float(lower, upper, shape, device):
    return (upper - lower) * torch.rand(*shape, device=device) + lower


def torch_rand_sqrt_float(lower, upper, shape, device):
    r = 2 * torch.rand(*shape, device=device) - 1
    r = torch.where(r < 0., -torch.sqrt(-r), torch.sqrt(r))
    return (upper - lower) * (r + 1.) / 2. + lower
