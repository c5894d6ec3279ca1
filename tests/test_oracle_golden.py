"""CPU tier: the oracle restatement against the reference's own recorded outputs (tests/golden, produced by
tests/golden/make_golden.py from /root/reference) and, where /root/reference exists, against the reference
run live.  Pins the oracle before the GPU tier trusts it."""
import pytest
import torch

import helpers
from legged_gym_b200.utils import scenarios
from oracle import reference_loader


@pytest.mark.parametrize("name", list(scenarios.SCENARIOS))
def test_oracle_matches_reference_golden(name):
    n_resets = helpers.replay_golden_oracle(name)
    assert n_resets >= 6      # every fixture exercises resets (time-outs + contact terminations)


@pytest.mark.reference
@pytest.mark.skipif(not reference_loader.reference_available(), reason="/root/reference only exists in the build container")
@pytest.mark.parametrize("name", ["go1_rough_full", "cassie_flat", "go1_flat_cmd_curriculum"])
def test_oracle_bit_identical_to_live_reference(name):
    from oracle import validate_against_reference as v
    assert v.run_scenario(name, num_envs=96, steps=6, seed=3, verbose=False) > 0


@pytest.mark.reference
@pytest.mark.skipif(not reference_loader.reference_available(), reason="/root/reference only exists in the build container")
def test_oracle_compute_torques_bit_identical_to_live_reference():
    """lo.compute_torques against LeggedRobot._compute_torques (legged_robot.py:370-392) for P / V / T."""
    from oracle import validate_against_reference as v
    assert v.run_compute_torques(num_envs=96, seed=1, verbose=False)


def test_command_curriculum_fixture_widens_the_range_on_its_gated_step():
    """update_command_curriculum (legged_robot.py:465-474) as the reference ran it: the recorded range steps
    from the shipped [-1, 1] to [-1.5, 1.5] on the third step of the fixture and nowhere else."""
    g = helpers.load_golden("go1_flat_cmd_curriculum")
    cr = g["out_command_ranges"][:, 0]
    assert (cr[:2] == cr[0]).all() and (cr[2:] == cr[2]).all()
    assert cr[2][0] == cr[0][0] - 0.5 and cr[2][1] == cr[0][1] + 0.5


@pytest.mark.reference
@pytest.mark.skipif(not reference_loader.reference_available(), reason="/root/reference only exists in the build container")
def test_sea_lstm_oracle_matches_the_shipped_torchscript_checkpoint():
    from oracle import validate_against_reference as v
    assert v.run_sea_lstm(num_envs=32, steps=12, verbose=False) < 1e-6


def test_golden_fixtures_cover_the_edge_cases():
    """resample on step 0, time-outs on steps 0 and 2, push on step 2, contact terminations somewhere."""
    g = helpers.load_golden("go1_rough_full")
    assert (g["out_resample_ids"][0] >= 0).sum() >= 8
    assert g["out_time_out_buf"][0].sum() >= 3 and g["out_time_out_buf"][2].sum() >= 2
    assert (g["out_reset_buf"] & ~g["out_time_out_buf"]).sum() >= 1
    # the push rewrites root lin vel xy of every env on the third step (common_step_counter % 750 == 0)
    assert not (g["out_root_states"][2][:, 7:9] == g["in_root_states"][2][:, 7:9]).all(axis=1).any()
    nores = ~g["out_reset_buf"][1]
    assert (g["out_root_states"][1][nores, 7:9] == g["in_root_states"][1][nores, 7:9]).all()
    # heights are int16 * 0.005f
    mh = g["out_measured_heights"][0]
    assert mh.shape == (64, 187) and (mh != 0).any()


# ---- the actuator-net checkpoints the reference ships (as fixture arrays), and what the reference computes with them
def _shipped():
    import numpy as np
    import os
    d = np.load(os.path.join(helpers.GOLDEN_DIR, "shipped_actuator_nets.npz"))
    return {k: d[k] for k in d.files}


def test_go1_actuator_oracle_on_the_shipped_checkpoint():
    """go1_net.pt (go1.py:45-47): ActuatorOracle with the shipped weights against the reference's own
    Go1.actuator_advance outputs, bit for bit (same torch CPU kernels, same order)."""
    from oracle import legged_oracle as lo
    g = _shipped()
    sd = {k[4:]: torch.from_numpy(v) for k, v in g.items() if k.startswith("go1.")}
    K, N = g["go1_in_actions"].shape[:2]
    act = lo.ActuatorOracle(N, sd)
    for k in range(K):
        ds = torch.from_numpy(g["go1_in_dof_state"][k])
        dv = act.advance(torch.from_numpy(g["go1_in_actions"][k]), ds[..., 0], ds[..., 1])
        assert torch.equal(dv, torch.from_numpy(g["go1_out_dvel"][k])), f"substep {k}"
    assert torch.equal(act.model_ins, torch.from_numpy(g["go1_out_model_ins"]))


def test_anymal_lstm_oracle_on_the_shipped_checkpoint():
    """anydrive_v3_lstm.pt (anymal.py:53-54): SeaLstmOracle's written-out cell, free-running, against the reference's
    Anymal._compute_torques (aten::lstm orders the gate sums differently: a few ulp on the state, 1e-6 of the
    output layer's operand magnitude on the torques)."""
    from oracle import legged_oracle as lo
    g = _shipped()
    sd = {k[7:]: torch.from_numpy(v) for k, v in g.items() if k.startswith("anymal.")}
    K, N = g["sea_in_actions"].shape[:2]
    sea = lo.SeaLstmOracle(N, 12, sd)
    env = helpers.make_env("anymal_c_flat", N, device="cpu")
    p = lo.OracleParams.from_env(env, device="cpu")
    atol = helpers.sea_atol_of(sd)
    for k in range(K):
        ds = torch.from_numpy(g["sea_in_dof_state"][k])
        tq = sea.compute_torques(torch.from_numpy(g["sea_in_actions"][k]), p.action_scale, p.default_dof_pos, ds[..., 0], ds[..., 1])
        helpers.assert_close_abs(f"sea torques[{k}]", tq, torch.from_numpy(g["sea_out_torques"][k]), atol)
        assert float((sea.sea_hidden_state - torch.from_numpy(g["sea_out_hidden"][k])).abs().max()) < 2e-6
        assert float((sea.sea_cell_state - torch.from_numpy(g["sea_out_cell"][k])).abs().max()) < 2e-6
