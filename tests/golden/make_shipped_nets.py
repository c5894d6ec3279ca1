"""Generates tests/golden/shipped_actuator_nets.npz: the two actuator-net checkpoints the reference ships
(resources/actuator_nets/go1_net.pt, envs/go1/go1.py:45-47; anydrive_v3_lstm.pt, envs/anymal_c/anymal.py:53-54)
as plain fp32 arrays, together with what the REFERENCE's own, unmodified task classes compute with them on seeded
inputs:

  go1.<key> / anymal.<key>     the checkpoints' state_dict entries
  go1_in_actions / go1_in_dof_state [K,N,12] / [K,N,12,2]   inputs of substep k
  go1_out_dvel [K,N,12], go1_out_model_ins [N,120]           Go1.actuator_advance (go1.py:79-107) per substep
  sea_in_actions / sea_in_dof_state                          inputs of call k
  sea_out_torques [K,N*12], sea_out_hidden / sea_out_cell [K,2,N*12,8]   Anymal._compute_torques (anymal.py:71-81),
                                                             free-running state

The checkpoints do not travel to the GPU box; these arrays do, so the CUDA path is checked on the shipped networks
there.  Only runs where /root/reference exists.

    python tests/golden/make_shipped_nets.py
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from legged_gym_b200.utils import scenarios, synthetic  # noqa: E402
from oracle.reference_loader import make_reference_env  # noqa: E402

K = 7      # more substeps than the Go1 history is deep
N = 48


def np_(t):
    return t.detach().cpu().numpy().copy()


def main():
    d = {}
    # ---- Go1: per-leg MLP over the 5-deep joint-error / velocity history --------------------------------------
    task, mutate = scenarios.SCENARIOS["go1_rough_actuator"]
    env = make_reference_env(task, mutate, num_envs=N, seed=0)
    for k, v in env.actuator_network.core_model.state_dict().items():
        d["go1." + k] = np_(v.float())
    frames = synthetic.make_frames(env, K, seed=4242)
    acts, dofs, dvels = [], [], []
    for fr in frames:
        synthetic.load_frame(env, fr)
        a = fr["actions"].clip(-env.cfg.normalization.clip_actions, env.cfg.normalization.clip_actions)
        acts.append(np_(a)); dofs.append(np_(env.dof_state.view(N, 12, 2)))
        dvels.append(np_(env.actuator_advance(a)))
    d["go1_in_actions"], d["go1_in_dof_state"], d["go1_out_dvel"] = np.stack(acts), np.stack(dofs), np.stack(dvels)
    d["go1_out_model_ins"] = np_(env.model_ins)
    # ---- ANYmal: LSTM series-elastic actuator, free-running ----------------------------------------------------
    task, mutate = scenarios.SCENARIOS["anymal_c_flat"]
    env = make_reference_env(task, mutate, num_envs=N, seed=0)
    for k, v in env.actuator_network.state_dict().items():
        d["anymal." + k] = np_(v.float())
    frames = synthetic.make_frames(env, K, seed=2424)
    acts, dofs, tqs, hid, cell = [], [], [], [], []
    for fr in frames:
        synthetic.load_frame(env, fr)
        a = fr["actions"].clip(-env.cfg.normalization.clip_actions, env.cfg.normalization.clip_actions)
        acts.append(np_(a)); dofs.append(np_(env.dof_state.view(N, 12, 2)))
        tqs.append(np_(env._compute_torques(a)).reshape(-1))
        hid.append(np_(env.sea_hidden_state)); cell.append(np_(env.sea_cell_state))
    d["sea_in_actions"], d["sea_in_dof_state"] = np.stack(acts), np.stack(dofs)
    d["sea_out_torques"], d["sea_out_hidden"], d["sea_out_cell"] = np.stack(tqs), np.stack(hid), np.stack(cell)
    path = os.path.join(HERE, "shipped_actuator_nets.npz")
    np.savez_compressed(path, **d)
    print(f"{path}: {os.path.getsize(path) / 1e3:.0f} kB, {len(d)} arrays")


if __name__ == "__main__":
    main()
