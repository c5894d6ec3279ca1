"""Go1 task: LeggedRobot + the per-leg actuator-net MLP over a 5-sample joint-error / velocity
history (reference ``envs/go1/go1.py:17-107``).

The reference keeps the histories as numpy float64 arrays on the host and moves 24 slices across
PCIe every simulator substep; here histories, packing and the 30-128-128-128-3 tanh MLP stay on the
GPU (``lgk_actuator_advance``).  As in the reference, the network output ``dVel`` is computed and
then not used by ``_compute_poses`` (go1.py:70-73); it is kept in ``self.dVel`` for inspection.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from legged_gym_b200 import LEGGED_GYM_ROOT_DIR, _lgk
from legged_gym_b200.envs.base.legged_robot import LeggedRobot

LEG_NUM = 4
LEG_DOF = 3
LEN_HIST = 5
MODEL_IN_SIZE = 2 * LEG_DOF * LEN_HIST

# normalisers of the joint error / velocity inputs, per (hip, thigh, calf) (go1.py:50-53)
POS_ERR_MEAN = (0.00036437, 0.01540757, -0.00972657)
POS_ERR_STD = (0.11722939, 0.19275887, 0.28700321)
VEL_MEAN = (-0.00017714, -0.00024455, 0.0005956)
VEL_STD = (2.31517027, 3.84613839, 5.52599008)


def pack_mlp_weights(state_dict) -> torch.Tensor:
    """fp32 blob W0[128,30] b0 W1[128,128] b1 W2[128,128] b2 W3[3,128] b3 in layer order."""
    keys = sorted(state_dict.keys(), key=lambda k: (int(k.split('.')[-2]), 0 if k.endswith('weight') else 1))
    return torch.cat([state_dict[k].detach().float().flatten().cpu() for k in keys])


def random_mlp_state_dict(seed: int = 0):
    """Same architecture as go1_net.pt with torch-default init (synthetic weights for tests / bench)."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for idx, (n_in, n_out) in zip((0, 2, 4, 6), ((MODEL_IN_SIZE, 128), (128, 128), (128, 128), (128, LEG_DOF))):
        bound = 1.0 / (n_in ** 0.5)
        sd[f"architecture.{idx}.weight"] = (torch.rand(n_out, n_in, generator=g) * 2 - 1) * bound
        sd[f"architecture.{idx}.bias"] = (torch.rand(n_out, generator=g) * 2 - 1) * bound
    return sd


class Go1(LeggedRobot):
    def __init__(self, cfg, sim_params, physics_engine, sim_device, headless):
        super().__init__(cfg, sim_params, physics_engine, sim_device, headless)
        self.pos_err_mean = torch.tensor(POS_ERR_MEAN).tile((LEG_NUM,)).to(self.device)
        self.pos_err_std = torch.tensor(POS_ERR_STD).tile((LEG_NUM,)).to(self.device)
        self.vel_mean = torch.tensor(VEL_MEAN).tile((LEG_NUM,)).to(self.device)
        self.vel_std = torch.tensor(VEL_STD).tile((LEG_NUM,)).to(self.device)
        self._act_norm = torch.cat([self.pos_err_mean, self.pos_err_std, self.vel_mean, self.vel_std]).contiguous()
        if getattr(self.cfg.control, "use_actuator_network", False):
            self.actuator_weights = self._load_actuator_weights().to(self.device)
            if self.device.startswith("cuda") if isinstance(self.device, str) else self.device.type == "cuda":
                # the tcgen05 MLP kernel's weight images live in memory this env owns (one workspace per weight set)
                nbytes = int(self._kernels().lgk_actuator_workspace_bytes())
                self._actuator_workspace = torch.zeros((nbytes + 7) // 8, dtype=torch.int64, device=self.device)
                # zero envs = set-up only: the kernel's host-side copy of the biases / output layer is read back
                # here, so that the first real call may already sit inside a CUDA-graph capture
                P = _lgk.ptr
                _lgk.check("lgk_actuator_advance", self._kernels().lgk_actuator_advance(
                    0, P(self.dVel), P(self.dof_state), P(self.pos_err_buffs), P(self.vel_buffs), P(self.model_ins),
                    P(self.actuator_weights), P(self._act_norm), P(self.dVel), P(self._actuator_workspace),
                    _lgk.current_stream(self.device)))

    def _load_actuator_weights(self) -> torch.Tensor:
        sd = getattr(self.cfg.control, "actuator_net_state_dict", None)
        if sd is not None:       # weights handed over in memory (e.g. the shipped checkpoint as fixture arrays)
            return pack_mlp_weights({k: torch.as_tensor(v) for k, v in sd.items()})
        seed = getattr(self.cfg.control, "actuator_net_random_seed", None)
        if seed is not None:
            return pack_mlp_weights(random_mlp_state_dict(int(seed)))
        path = self.cfg.control.actuator_net_file.format(LEGGED_GYM_ROOT_DIR=LEGGED_GYM_ROOT_DIR)
        if not os.path.isfile(path) and os.environ.get("LEGGED_GYM_RESOURCES"):
            # checkpoints are not shipped with this package; point this at <reference>/resources
            path = self.cfg.control.actuator_net_file.format(
                LEGGED_GYM_ROOT_DIR=os.path.dirname(os.environ["LEGGED_GYM_RESOURCES"].rstrip("/")))
        if not os.path.isfile(path):
            raise FileNotFoundError(
                f"actuator net checkpoint not found: {path} (set cfg.control.actuator_net_file, or "
                f"cfg.control.actuator_net_random_seed for synthetic weights)")
        module = torch.jit.load(path, map_location="cpu")
        return pack_mlp_weights(module.state_dict())

    def _init_buffers(self):
        super()._init_buffers()
        kw = dict(device=self.device, dtype=torch.float, requires_grad=False)
        self.model_ins = torch.zeros(self.num_envs, MODEL_IN_SIZE * LEG_NUM, **kw)
        # fp32 on the device (the reference's float64 host arrays hold fp32 values exactly, go1.py:65-66)
        self.pos_err_buffs = torch.zeros(self.num_envs, self.num_actions, LEN_HIST, **kw)
        self.vel_buffs = torch.zeros(self.num_envs, self.num_actions, LEN_HIST, **kw)
        self.dVel = torch.zeros(self.num_envs, self.num_actions, **kw)

    def _extra_state_tensors(self):
        return {"pos_err_buffs": self.pos_err_buffs, "vel_buffs": self.vel_buffs, "model_ins": self.model_ins}

    def _compute_poses(self, actions):
        if getattr(self.cfg.control, "use_actuator_network", False):
            self.actuator_advance(actions)   # result unused, as in the reference (go1.py:70-73)
        return super()._compute_poses(actions)

    def actuator_advance(self, actions):
        """History shift + per-leg MLP; returns dVel [N,12] (go1.py:79-107)."""
        lib = self._kernels()
        P = _lgk.ptr
        _lgk.check("lgk_actuator_advance", lib.lgk_actuator_advance(
            self.num_envs, P(actions), P(self.dof_state), P(self.pos_err_buffs), P(self.vel_buffs),
            P(self.model_ins), P(self.actuator_weights), P(self._act_norm), P(self.dVel),
            P(getattr(self, "_actuator_workspace", None)), _lgk.current_stream(self.device)))
        return self.dVel
