"""Restatement of the ``isaacgym.torch_utils`` functions on the hot path -- TEST INFRASTRUCTURE.

Isaac Gym is a closed, un-vendored dependency of the reference (``setup.py:12``, no version pin;
README asks for Preview 3).  Its python ``torch_utils`` source is not under /root/reference; the
six functions below follow the published Preview-4 ``python/isaacgym/torch_utils.py`` algorithm as
recorded in SURVEY.md Appendix A-1 (Hamilton quaternions, xyzw).  Call sites in the reference:
``legged_robot.py:123-125,345,530-531,548-550``, ``utils/math.py:41-42``.  Parity at this boundary is
"unpinned": the reference holds no test vectors for these functions.
"""
import numpy as np
import torch


def to_torch(x, dtype=torch.float, device='cuda:0', requires_grad=False):
    return torch.tensor(x, dtype=dtype, device=device, requires_grad=requires_grad)


def get_axis_params(value, axis_idx, x_value=0., dtype=float, n_dims=3):
    zs = np.zeros((n_dims,))
    assert axis_idx < n_dims
    zs[axis_idx] = 1.
    params = np.where(zs == 1., value, zs)
    params[0] = x_value
    return list(params.astype(dtype))


def normalize(x, eps: float = 1e-9):
    return x / x.norm(p=2, dim=-1).clamp(min=eps, max=None).unsqueeze(-1)


def quat_apply(a, b):
    shape = b.shape
    a = a.reshape(-1, 4)
    b = b.reshape(-1, 3)
    xyz = a[:, :3]
    t = xyz.cross(b, dim=-1) * 2
    return (b + a[:, 3:] * t + xyz.cross(t, dim=-1)).view(shape)


def quat_rotate_inverse(q, v):
    shape = q.shape
    q_w = q[:, -1]
    q_vec = q[:, :3]
    a = v * (2.0 * q_w ** 2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * torch.bmm(q_vec.view(shape[0], 1, 3), v.view(shape[0], 3, 1)).squeeze(-1) * 2.0
    return a - b + c


def torch_rand_float(lower, upper, shape, device):
    return (upper - lower) * torch.rand(*shape, device=device) + lower
