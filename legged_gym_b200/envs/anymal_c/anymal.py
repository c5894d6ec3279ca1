"""ANYmal task: LeggedRobot + the series-elastic-actuator LSTM behind ``_compute_torques``
(reference ``envs/anymal_c/anymal.py:46-81``, checkpoint ``resources/actuator_nets/anydrive_v3_lstm.pt``).

Per joint the network is a 2-layer ``torch.nn.LSTM(2, 8)`` stepped once per simulator substep plus a
``Linear(8, 1)``; the reference calls the TorchScript module on ``[N*D, 1, 2]`` inputs (cuDNN RNN + four
small kernels).  Here ``lgk_lstm_actuator_advance`` does the input packing, both cells and the output
layer in one launch, one thread per joint, with the hidden / cell state updated in place in the same
``sea_hidden_state`` / ``sea_cell_state`` tensors the reference exposes.  Resets (``anymal.py:56-60``) zero
the state of the reset envs from the step kernel's device-side id list, without a host sync.

As in the reference, this fork's ``step()`` drives the joints with position targets and never calls
``_compute_torques`` (``legged_robot.py:90-95``); a caller in torque mode gets the LSTM torques from it.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from legged_gym_b200 import LEGGED_GYM_ROOT_DIR, _lgk
from legged_gym_b200.envs.base.legged_robot import LeggedRobot

SEA_HIDDEN = _lgk.LGK_LSTM_HID
SEA_LAYERS = 2

_LSTM_KEYS = ("lstm.weight_ih_l0", "lstm.weight_hh_l0", "lstm.bias_ih_l0", "lstm.bias_hh_l0",
              "lstm.weight_ih_l1", "lstm.weight_hh_l1", "lstm.bias_ih_l1", "lstm.bias_hh_l1",
              "linear.weight", "linear.bias", "in_scale", "out_scale")
_LSTM_FIELDS = ("w_ih0", "w_hh0", "b_ih0", "b_hh0", "w_ih1", "w_hh1", "b_ih1", "b_hh1", "lin_w", "lin_b",
                "in_scale", "out_scale")


def random_lstm_state_dict(seed: int = 0):
    """Same architecture as anydrive_v3_lstm.pt (LSTMsea: in_scale, LSTM(2, 8, 2 layers), Linear(8, 1),
    out_scale) with torch-default uniform(-1/sqrt(8), 1/sqrt(8)) init -- synthetic weights for tests / bench."""
    g = torch.Generator().manual_seed(seed)
    k = 1.0 / (SEA_HIDDEN ** 0.5)
    u = lambda *shape: (torch.rand(*shape, generator=g) * 2 - 1) * k
    G = 4 * SEA_HIDDEN
    return {
        "lstm.weight_ih_l0": u(G, 2), "lstm.weight_hh_l0": u(G, SEA_HIDDEN), "lstm.bias_ih_l0": u(G),
        "lstm.bias_hh_l0": u(G), "lstm.weight_ih_l1": u(G, SEA_HIDDEN), "lstm.weight_hh_l1": u(G, SEA_HIDDEN),
        "lstm.bias_ih_l1": u(G), "lstm.bias_hh_l1": u(G), "linear.weight": u(1, SEA_HIDDEN), "linear.bias": u(1),
        "in_scale": torch.tensor([[[2.5, 0.08]]]), "out_scale": torch.tensor([40.0]),
    }


def pack_lstm_weights(state_dict) -> "_lgk.LgkLstmWeights":
    """state_dict of the reference's LSTMsea module -> the C struct (host memory, include/lgk.h)."""
    w = _lgk.LgkLstmWeights()
    for key, field in zip(_LSTM_KEYS, _LSTM_FIELDS):
        flat = state_dict[key].detach().float().flatten().cpu().tolist()
        if field in ("lin_b", "out_scale"):
            setattr(w, field, flat[0])
        else:
            arr = getattr(w, field)
            assert len(flat) == len(arr), f"{key}: {len(flat)} values, expected {len(arr)}"
            for i, v in enumerate(flat):
                arr[i] = v
    return w


class Anymal(LeggedRobot):
    def __init__(self, cfg, sim_params, physics_engine, sim_device, headless):
        super().__init__(cfg, sim_params, physics_engine, sim_device, headless)
        # load actuator network (anymal.py:51-54)
        self._sea_weights = None
        if self.cfg.control.use_actuator_network:
            self.actuator_state_dict = self._load_actuator_state_dict()
            self._sea_weights = pack_lstm_weights(self.actuator_state_dict)

    def _load_actuator_state_dict(self):
        sd = getattr(self.cfg.control, "actuator_net_state_dict", None)
        if sd is not None:       # weights handed over in memory (e.g. the shipped checkpoint as fixture arrays)
            return {k: torch.as_tensor(v) for k, v in sd.items()}
        seed = getattr(self.cfg.control, "actuator_net_random_seed", None)
        if seed is not None:
            return random_lstm_state_dict(int(seed))
        path = self.cfg.control.actuator_net_file.format(LEGGED_GYM_ROOT_DIR=LEGGED_GYM_ROOT_DIR)
        if not os.path.isfile(path) and os.environ.get("LEGGED_GYM_RESOURCES"):
            # checkpoints are not shipped with this package; point this at <reference>/resources
            path = self.cfg.control.actuator_net_file.format(
                LEGGED_GYM_ROOT_DIR=os.path.dirname(os.environ["LEGGED_GYM_RESOURCES"].rstrip("/")))
        if not os.path.isfile(path):
            raise FileNotFoundError(
                f"actuator net checkpoint not found: {path} (set cfg.control.actuator_net_file, or "
                f"cfg.control.actuator_net_random_seed for synthetic weights)")
        return dict(torch.jit.load(path, map_location="cpu").state_dict())

    def reset_idx(self, env_ids):
        super().reset_idx(env_ids)
        if len(env_ids) == 0:
            return
        # Additionaly empty actuator network hidden states (anymal.py:58-60)
        ids = env_ids.to(device=self.device, dtype=torch.int32).contiguous()
        _lgk.check("lgk_lstm_reset_hidden", self._kernels().lgk_lstm_reset_hidden(
            self.num_envs, self.num_actions, _lgk.ptr(ids), None, int(len(ids)), _lgk.ptr(self.sea_hidden_state),
            _lgk.ptr(self.sea_cell_state), _lgk.current_stream(self.device)))

    def _after_fused_resets(self):
        """The per-step resets happen inside the fused kernel; clear the LSTM state of this step's
        reset envs from its device-side id list and count (no host sync)."""
        _lgk.check("lgk_lstm_reset_hidden", self._kernels().lgk_lstm_reset_hidden(
            self.num_envs, self.num_actions, _lgk.ptr(self.reset_ids), _lgk.ptr(self.step_counts), 0,
            _lgk.ptr(self.sea_hidden_state), _lgk.ptr(self.sea_cell_state), _lgk.current_stream(self.device)))

    def _init_buffers(self):
        super()._init_buffers()
        # Additionally initialize actuator network hidden state tensors (anymal.py:64-69)
        kw = dict(device=self.device, dtype=torch.float, requires_grad=False)
        M = self.num_envs * self.num_actions
        self.sea_input = torch.zeros(M, 1, 2, **kw)
        self.sea_hidden_state = torch.zeros(SEA_LAYERS, M, SEA_HIDDEN, **kw)
        self.sea_cell_state = torch.zeros(SEA_LAYERS, M, SEA_HIDDEN, **kw)
        self.sea_hidden_state_per_env = self.sea_hidden_state.view(SEA_LAYERS, self.num_envs, self.num_actions, SEA_HIDDEN)
        self.sea_cell_state_per_env = self.sea_cell_state.view(SEA_LAYERS, self.num_envs, self.num_actions, SEA_HIDDEN)
        self._sea_torques = torch.zeros(M, **kw)

    def _extra_state_tensors(self):
        return {"sea_hidden_state": self.sea_hidden_state, "sea_cell_state": self.sea_cell_state}

    def _compute_torques(self, actions):
        # Choose between pd controller and actuator network (anymal.py:71-81)
        if self.cfg.control.use_actuator_network:
            lib = self._kernels()
            P = _lgk.ptr
            actions = actions.to(self.device, dtype=torch.float).contiguous()
            _lgk.check("lgk_lstm_actuator_advance", lib.lgk_lstm_actuator_advance(
                C.byref(self._consts), C.byref(self._sea_weights), P(actions), P(self.dof_state), P(self.sea_input),
                P(self.sea_hidden_state), P(self.sea_cell_state), P(self._sea_torques),
                _lgk.current_stream(self.device)))
            return self._sea_torques          # flat [N*D], like the reference's squeeze()d network output
        # pd controller
        return super()._compute_torques(actions)
