"""GPU tier: the CUDA path, called through the C ABI (liblgk.so), against
  (1) the reference's own recorded outputs (tests/golden),
  (2) the oracle restatement on identical frames + RNG taps (torch-CPU and torch-CUDA variants), and
  (3) size-independent properties at the full BASELINE.json size (65536 envs).
Tolerances (north star): bit-exact for terminations, id lists, height-field indices / measured heights,
episode lengths, terrain levels; |a-b| <= 1e-5*max(|a|,|b|) + 1e-6 for fp32 outputs (helpers.RTOL/ATOL)."""
import ctypes as C

import pytest
import torch

import helpers
from legged_gym_b200 import _lgk
from legged_gym_b200.utils import scenarios, synthetic

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module", autouse=True)
def _need_cuda():
    assert torch.cuda.is_available(), "GPU tier needs a CUDA device"
    _lgk.load()


# ------------------------------------------------------------------------------------------- (1) golden
@pytest.mark.parametrize("name", list(scenarios.SCENARIOS))
def test_cuda_path_matches_reference_golden(name):
    assert helpers.replay_golden_env(name) >= 6


# ------------------------------------------------------------------------------------------- (2) oracle
@pytest.mark.parametrize("oracle_device", ["cpu", "cuda:0"])
@pytest.mark.parametrize("name", list(scenarios.SCENARIOS))
def test_step_matches_oracle(name, oracle_device):
    n = helpers.run_parity(name, num_envs=1024, steps=6, oracle_device=oracle_device)
    assert n > 0


@pytest.mark.parametrize("name,envs", [("go1_flat", 1800), ("go1_rough_full", 4096), ("go1_rough_actuator", 16384),
                                       ("cassie_flat", 8192)])
def test_step_matches_oracle_at_the_baseline_sizes(name, envs):
    """BASELINE.json configs 1-4 at their own env counts: these sizes take other grid shapes than 1 024 and 65 536
    (single-wave small tiles of the step kernel, shorter scan tiles)."""
    assert helpers.run_parity(name, num_envs=envs, steps=3, seed=2) > 0


@pytest.mark.parametrize("name", list(scenarios.SHIPPED_TASKS))
def test_every_registered_task_as_shipped(name):
    """The eight registrations of the reference (envs/__init__.py:52-59), configs untouched."""
    assert helpers.run_parity(name, 200, 3) >= 0


def test_blind_flat_task_with_every_reward_term():
    """Flat, blind Go1 (no height scan: measured_heights stays the python int 0, legged_robot.py:553) with all 20
    terms + termination: base_height then reduces to (root_z - target)^2 (:869-872)."""
    def full(cfg):
        for k, v in scenarios.FULL_REWARD_EXTRA.items():
            setattr(cfg.rewards.scales, k, v)
    assert helpers.run_parity("go1_flat", 515, 5, extra=full) > 0
    assert helpers.run_parity("go1_flat", 515, 3, extra=full, oracle_device="cuda:0") > 0


@pytest.mark.parametrize("scenario", ["go1_rough_full", "a1_variant"])
def test_device_terrain_generation_matches_the_host_generator(scenario):
    """cfg.terrain.device_generation: lgk_generate_terrain fills the int16 height field and
    the env-origin heights from the host's recipes (same np.random stream) -- bit-identical to the host generators for
    the curriculum grid (go1_rough_full: every sub-terrain type of make_terrain) and for randomised sub-terrains
    (a1_variant), and therefore to the reference's field the golden fixtures were recorded on."""
    def dev(cfg):
        cfg.terrain.device_generation = True
    host = helpers.make_env(scenario, 64, device=DEV, seed=0)
    devi = helpers.make_env(scenario, 64, device=DEV, seed=0, extra=dev)
    assert devi.terrain.height_field_dev is not None and devi.height_samples.is_cuda
    assert devi.height_samples.data_ptr() == devi.terrain.height_field_dev.data_ptr()      # no host round trip
    assert torch.equal(host.height_samples, devi.height_samples)
    assert torch.equal(host.terrain_origins, devi.terrain_origins)
    assert torch.equal(host.env_origins, devi.env_origins)
    assert (host.terrain.height_field_raw == devi.terrain.height_field_raw).all()
    g = helpers.load_golden(scenario)
    assert helpers._crc(devi.height_samples) == int(g["meta_terrain_crc"])
    assert helpers._crc(devi.terrain_origins) == int(g["meta_origins_crc"])
    # and a step on it
    assert helpers.run_parity(scenario, 256, 2, extra=dev) >= 0


def test_filtered_feet_contact_variant_of_feet_air_time():
    """cfg.rewards.filtered_feet_contact: _reward_feet_air_time as the reference keeps it commented out
    (legged_robot.py:928-939) -- a foot is in contact if it is now or was at the previous step, last_contacts is real
    state.  Against the oracle's restatement of those lines (no executable reference: parity unpinned for this option);
    the results must also differ from the unfiltered term."""
    def filt(cfg):
        cfg.rewards.filtered_feet_contact = True
    assert helpers.run_parity("go1_rough_full", 777, 6, extra=filt) > 0
    assert helpers.run_parity("cassie_flat", 300, 4, extra=filt, oracle_device="cuda:0") >= 0
    # and it is not a no-op: same frames with and without the filter give different air-time state
    envs = []
    for extra in (filt, None):
        env = helpers.make_env("go1_rough_full", 256, device=DEV, extra=extra)
        helpers.randomise_env_state(env, 0)
        for it, fr in enumerate(synthetic.make_frames(env, 3, seed=5)):
            synthetic.load_frame(env, fr)
            env.use_rng_taps(synthetic.make_taps(env, seed=it))
            env.step(fr["actions"].clone())
        torch.cuda.synchronize()
        envs.append(env)
    assert envs[0].last_contacts.any() and not envs[1].last_contacts.any()
    assert not torch.equal(envs[0].feet_air_time, envs[1].feet_air_time)


def test_cassie_on_rough_terrain_scans_121_points():
    """Cassie as shipped sits on a plane (121 zero heights); on a trimesh the 4-pass scan instantiation reads real
    terrain, with two feet, the pelvis termination bodies and the terrain curriculum."""
    def rough(cfg):
        cfg.terrain.mesh_type = 'trimesh'
        cfg.terrain.curriculum = True
    assert helpers.run_parity("cassie_flat", 300, 4, extra=rough) > 0


def test_generic_scan_width_256_points():
    """A scan grid no shipped task uses (16 x 16 = 256 points, 304 observations): runs on the generic (not
    specialised) instantiation of the step kernel with 8 passes."""
    def wide(cfg):
        scenarios.go1_rough_full(cfg)
        cfg.terrain.measured_points_x = [round(-0.75 + 0.1 * i, 2) for i in range(16)]
        cfg.terrain.measured_points_y = [round(-0.75 + 0.1 * i, 2) for i in range(16)]
        cfg.env.num_observations = 48 + 256
    assert helpers.run_parity("go1_flat", 200, 3, extra=wide) > 0


def test_scan_of_357_points_runs_in_two_chunks():
    """21 x 17 = 357 scan points (more than the 256 the generic scan instantiation takes per chunk; round 1
    rejected such a grid with LGK_ERR_DIM): two chunks per env, two Philox blocks per lane."""
    def wide(cfg):
        scenarios.go1_rough_full(cfg)
        cfg.terrain.measured_points_x = [round(-1.0 + 0.1 * i, 2) for i in range(21)]
        cfg.terrain.measured_points_y = [round(-0.8 + 0.1 * i, 2) for i in range(17)]
        cfg.env.num_observations = 48 + 357
    assert helpers.run_parity("go1_flat", 200, 3, extra=wide) > 0
    assert helpers.run_parity("go1_flat", 130, 2, extra=wide, oracle_device="cuda:0") > 0


def _ten_dof_biped(cfg):
    """A robot no shipped task has: 10 joints (not a multiple of 4: the step kernel's scalar joint path, no bulk
    copies of the joint rows), 11 bodies, two feet."""
    from legged_gym_b200.sim import assets
    names = [f"{side}_{j}_joint" for side in ("left", "right") for j in ("hip_roll", "hip_pitch", "knee", "ankle", "toe")]
    bodies = ["pelvis"] + [f"{side}_{b}" for side in ("left", "right") for b in ("hip", "thigh", "shin", "tarsus", "toe")]
    assets.BUILTIN_ASSETS["biped10"] = assets.RobotAsset(
        "biped10", names, bodies, lower=[-0.5 - 0.05 * i for i in range(10)], upper=[0.6 + 0.05 * i for i in range(10)],
        velocity=[8.0 + i for i in range(10)], effort=[40.0 + 3 * i for i in range(10)], body_mass=[1.0] * 11)
    scenarios.go1_rough_full(cfg)
    cfg.asset.file = "{LEGGED_GYM_ROOT_DIR}/resources/robots/biped10/urdf/biped10.urdf"
    cfg.asset.name = "biped10"
    cfg.asset.foot_name = "toe"
    cfg.asset.penalize_contacts_on = ["shin", "tarsus"]
    cfg.asset.terminate_after_contacts_on = ["pelvis"]
    cfg.init_state.default_joint_angles = {n: 0.05 * (i - 4) for i, n in enumerate(names)}
    cfg.control.stiffness = {"hip_roll": 100., "hip_pitch": 120., "knee": 200., "ankle": 60., "toe": 25.}
    cfg.control.damping = {"hip_roll": 3., "hip_pitch": 3.5, "knee": 6., "ankle": 2., "toe": 1.}
    cfg.env.num_actions = 10
    cfg.env.num_observations = 12 + 3 * 10 + 187


def test_ten_dof_robot_takes_the_generic_instantiation():
    """D = 10: the <DT = 0> step kernel (scalar joint loop, Philox block per joint pair rotated into place), the
    finish kernel and the scan kernel for a body / joint layout none of the shipped robots has, against the oracle
    (rough terrain, all reward terms, both oracle devices), plus the substep kernels' scalar paths."""
    assert helpers.run_parity("a1_variant", 517, 4, extra=_ten_dof_biped) > 0
    assert helpers.run_parity("a1_variant", 96, 3, extra=_ten_dof_biped, oracle_device="cuda:0") >= 0


@pytest.mark.parametrize("num_envs", [1, 3, 33, 127, 1800])
def test_ragged_env_counts(num_envs):
    """env counts that are not multiples of the warp / CTA tile (config 1 uses 1800)."""
    helpers.run_parity("go1_rough_full", num_envs=num_envs, steps=4, oracle_device="cpu", seed=2)
    helpers.run_parity("go1_flat", num_envs=num_envs, steps=3, oracle_device="cpu", seed=3)


@pytest.mark.parametrize("clip", [0.75, 5.2])
def test_tight_observation_clip(clip):
    """clip_observations inside the range of the scan columns (5 * 1 + noise): the kernel's copy of the
    scan loop that keeps the final clip runs (the default 100 lets it drop the clip)."""
    def tight(cfg):
        cfg.normalization.clip_observations = clip
    helpers.run_parity("go1_rough_full", num_envs=300, steps=3, oracle_device="cpu", seed=5, extra=tight)


@pytest.mark.parametrize("per_cta", [4, 40, 76, 200])
def test_persistent_cta_ranges(per_cta):
    """envs per persistent CTA: one partial sub-tile, 32+8, 32+32+12 (the 65536-env shape), many sub-tiles."""
    def force(cfg):
        cfg.env.envs_per_cta = per_cta
    env = helpers.make_env("go1_rough_full", 1000, device=DEV, seed=6, extra=force)
    ref = helpers.make_env("go1_rough_full", 1000, device=DEV, seed=6)
    fr = synthetic.make_frames(env, 2, seed=13)
    for it in range(3):
        taps = synthetic.make_taps(env, seed=40 + it)
        for e in (env, ref):
            if it == 0:
                helpers.randomise_env_state(e, 6)
            synthetic.load_frame(e, fr[it % 2])
            e.use_rng_taps(taps)
            e.step(fr[it % 2]["actions"].clone())
        torch.cuda.synchronize()
        for name in ("obs_buf", "rew_buf", "reset_buf", "episode_length_buf", "root_states", "dof_state", "commands",
                     "last_dof_vel", "last_actions", "feet_air_time", "terrain_levels", "measured_heights"):
            assert torch.equal(getattr(env, name), getattr(ref, name)), f"{name} differs at per_cta={per_cta}, step {it}"
        n = int(ref.step_counts[0])
        assert int(env.step_counts[0]) == n and torch.equal(env.reset_ids[:n], ref.reset_ids[:n])
        for k in env.episode_sums:
            assert torch.equal(env.episode_sums[k], ref.episode_sums[k])
        for k, v in ref.extras["episode"].items():
            if torch.is_tensor(v):
                helpers.assert_close(f"extras[{k}]", env.extras["episode"][k].reshape(1), v.reshape(1))


def test_heights_not_materialised_gives_same_observations():
    def nomat(cfg):
        cfg.terrain.materialize_heights = False
    a = helpers.make_env("go1_rough_full", 512, device=DEV, seed=5)
    b = helpers.make_env("go1_rough_full", 512, device=DEV, seed=5, extra=nomat)
    assert torch.is_tensor(a.measured_heights) and not torch.is_tensor(b.measured_heights)
    fr = synthetic.make_frames(a, 1, seed=11)[0]
    taps = synthetic.make_taps(a, seed=4)
    for env in (a, b):
        helpers.randomise_env_state(env, 5)
        synthetic.load_frame(env, fr)
        env.use_rng_taps(taps)
        env.step(fr["actions"].clone())
    assert torch.equal(a.obs_buf, b.obs_buf) and torch.equal(a.rew_buf, b.rew_buf)


def test_robots_outside_the_terrain_are_clipped_like_the_reference():
    from oracle import legged_oracle as lo
    env = helpers.make_env("go1_rough_full", 256, device=DEV, seed=1)
    helpers.randomise_env_state(env, 1)
    fr = synthetic.make_frames(env, 1, seed=5)[0]
    fr["root_states"][:, 0] = torch.linspace(-80., 220., 256, device=DEV)
    fr["root_states"][:, 1] = torch.linspace(330., -90., 256, device=DEV)
    synthetic.load_frame(env, fr)
    taps = synthetic.make_taps(env, seed=2)
    p = lo.OracleParams.from_env(env, device=DEV)
    s = lo.OracleState.from_env(env, p, device=DEV)
    env.use_rng_taps(taps, index_math=_lgk.INDEX_MATH_TORCH_CUDA)
    env.step(fr["actions"].clone())
    lo.step(p, s, fr["actions"].clone(), taps)
    helpers.compare_env_to_oracle(env, s, tag="[outside] ")


# ------------------------------------------------------------------------------------------- control
@pytest.mark.parametrize("control_type", ["P", "V", "T"])
def test_compute_torques(control_type):
    from oracle import legged_oracle as lo

    def ct(cfg):
        cfg.control.control_type = control_type
    env = helpers.make_env("go1_flat", 777, device=DEV, extra=ct)
    fr = synthetic.make_frames(env, 1, seed=3)[0]
    synthetic.load_frame(env, fr)
    env.last_dof_vel[:] = torch.randn_like(env.last_dof_vel)
    p = lo.OracleParams.from_env(env, device="cpu")
    s = lo.OracleState.from_env(env, p, device="cpu")
    a = (3 * torch.randn(777, 12, device=DEV))
    got = env._compute_torques(a)
    want = lo.compute_torques(p, s, a.cpu())
    helpers.assert_close(f"torques[{control_type}]", got, want)
    assert (got.abs() <= env.torque_limits + 1e-6).all()


def test_unknown_control_type_raises_name_error():
    env = helpers.make_env("go1_flat", 8, device=DEV)
    env.cfg.control.control_type = "Q"
    with pytest.raises(NameError):
        env._compute_torques(torch.zeros(8, 12, device=DEV))


@pytest.mark.parametrize("n", [1, 5, 4096])
def test_clip_actions_and_poses(n):
    from oracle import legged_oracle as lo
    env = helpers.make_env("cassie_flat", n, device=DEV)
    lib = _lgk.load()
    a = 200 * torch.randn(n, 12, device=DEV)
    out = torch.empty_like(a)
    _lgk.check("clip", lib.lgk_clip_actions(_lgk.ptr(a), _lgk.ptr(out), a.numel(), 100.0, _lgk.current_stream(DEV)))
    assert torch.equal(out, a.clip(-100, 100))
    # unaligned views take the scalar kernel
    flat = torch.empty(n * 12 + 1, device=DEV)
    _lgk.check("clip", lib.lgk_clip_actions(_lgk.ptr(a), C.c_void_p(flat.data_ptr() + 4), a.numel(), 1.0,
                                            _lgk.current_stream(DEV)))
    assert torch.equal(flat[1:].view(n, 12), a.clip(-1, 1))
    p = lo.OracleParams.from_env(env, device="cpu")
    helpers.assert_close("target_poses", env._compute_poses(out), lo.compute_poses(p, out.cpu()))


@pytest.mark.parametrize("kernel", ["umma", "mma", "ffma"])
def test_actuator_net_matches_oracle_over_substeps(kernel, monkeypatch):
    """All three MLP kernels: 3xTF32 on tcgen05 with the activations in tensor memory and 3xTF32 on mma.sync (floor
    of a few fp32 roundings of the output layer's operands, helpers.actuator_dvel_atol), and fp32 FFMA (plain
    1e-5 + 1e-6)."""
    from legged_gym_b200.envs.go1.go1 import random_mlp_state_dict
    monkeypatch.delenv("LGK_MLP_FFMA", raising=False)
    monkeypatch.setenv("LGK_MLP_KERNEL", kernel)
    from oracle import legged_oracle as lo
    n = 1000
    env = helpers.make_env("go1_rough_actuator", n, device=DEV)
    act = lo.ActuatorOracle(n, random_mlp_state_dict(0))
    frames = synthetic.make_frames(env, 7, seed=8)
    for it, fr in enumerate(frames):       # > LEN_HIST substeps so the whole window is exercised
        synthetic.load_frame(env, fr)
        a = fr["actions"]
        got = env.actuator_advance(a)
        want = act.advance(a.cpu(), env.dof_pos.cpu(), env.dof_vel.cpu())
        if kernel == "ffma":
            helpers.assert_close(f"dVel[{it}]", got, want)
        else:
            helpers.assert_close_abs(f"dVel[{it}]", got, want, helpers.actuator_dvel_atol(random_mlp_state_dict(0)))
        helpers.assert_close(f"model_ins[{it}]", env.model_ins, act.model_ins)
        assert torch.equal(env.pos_err_buffs.cpu().double(), torch.from_numpy(act.pos_err_buffs)) or \
            torch.allclose(env.pos_err_buffs.cpu().double(), torch.from_numpy(act.pos_err_buffs), rtol=1e-6, atol=1e-7)


def _shipped_nets():
    import os
    import numpy as np
    d = np.load(os.path.join(helpers.GOLDEN_DIR, "shipped_actuator_nets.npz"))
    return {k: d[k] for k in d.files}


@pytest.mark.parametrize("kernel", ["umma", "mma", "ffma"])
def test_go1_actuator_kernels_on_the_shipped_checkpoint(kernel, monkeypatch):
    """The network the reference ships (go1_net.pt, go1.py:45-47; its weights travel as fixture arrays) through the
    three MLP kernels, against what the reference's own Go1.actuator_advance computed with it
    (tests/golden/make_shipped_nets.py), over more substeps than the history is deep."""
    monkeypatch.delenv("LGK_MLP_FFMA", raising=False)
    monkeypatch.setenv("LGK_MLP_KERNEL", kernel)
    g = _shipped_nets()
    sd = {k[4:]: torch.from_numpy(v) for k, v in g.items() if k.startswith("go1.")}
    K, N = g["go1_in_actions"].shape[:2]

    def shipped(cfg):
        cfg.control.actuator_net_state_dict = sd
    env = helpers.make_env("go1_rough_actuator", N, device=DEV, extra=shipped)
    atol = helpers.actuator_dvel_atol(sd)
    outside = []
    for k in range(K):
        env.dof_state.copy_(torch.from_numpy(g["go1_in_dof_state"][k]).to(DEV).view_as(env.dof_state))
        got = env.actuator_advance(torch.from_numpy(g["go1_in_actions"][k]).to(DEV))
        want = torch.from_numpy(g["go1_out_dvel"][k])
        # The shipped output layer has larger weights than the synthetic one: dVel is a cancelling 128-term sum whose
        # fp32 evaluation error scales with its OPERANDS, so even the fp32 FFMA kernel (same arithmetic as torch, other
        # summation order) leaves a few elements a fraction of 1e-6 outside the plain 1e-5 rel + 1e-6 (DESIGN.md
        # section 5 records the fractions).  Bar here: the operand-scaled floor for all three kernels, and at most
        # 2 % of the elements outside the plain tolerance.
        helpers.assert_close_abs(f"dVel[{k}]", got, want, atol)
        a, b = got.cpu().double(), want.double()
        outside.append(float(((a - b).abs() > helpers.RTOL * torch.maximum(a.abs(), b.abs()) + helpers.ATOL).double().mean()))
    print(f"[shipped go1_net.pt, {kernel}] fraction of dVel elements outside plain 1e-5 rel + 1e-6 per substep: "
          + " ".join(f"{x:.4f}" for x in outside))
    # fp32 FFMA: a handful of elements; the 3xTF32 tensor-core kernels: up to ~11 % on the first substeps (histories
    # still mostly zero, small outputs), all within the operand-scaled floor asserted above
    assert max(outside) <= (0.02 if kernel == "ffma" else 0.15)
    helpers.assert_close("model_ins", env.model_ins, torch.from_numpy(g["go1_out_model_ins"]))


def test_anymal_lstm_kernel_on_the_shipped_checkpoint():
    """anydrive_v3_lstm.pt (anymal.py:53-54) as fixture arrays through lgk_lstm_actuator_advance, free-running,
    against the reference's own Anymal._compute_torques with the TorchScript module."""
    g = _shipped_nets()
    sd = {k[7:]: torch.from_numpy(v) for k, v in g.items() if k.startswith("anymal.")}
    K, N = g["sea_in_actions"].shape[:2]

    def shipped(cfg):
        cfg.control.actuator_net_state_dict = sd
    env = helpers.make_env("anymal_c_flat", N, device=DEV, extra=shipped)
    atol = helpers.sea_atol_of(sd)
    for k in range(K):
        env.dof_state.copy_(torch.from_numpy(g["sea_in_dof_state"][k]).to(DEV).view_as(env.dof_state))
        tq = env._compute_torques(torch.from_numpy(g["sea_in_actions"][k]).to(DEV))
        helpers.assert_close_abs(f"sea torques[{k}]", tq, torch.from_numpy(g["sea_out_torques"][k]), atol)
        assert float((env.sea_hidden_state.cpu() - torch.from_numpy(g["sea_out_hidden"][k])).abs().max()) < 5e-6
        assert float((env.sea_cell_state.cpu() - torch.from_numpy(g["sea_out_cell"][k])).abs().max()) < 5e-6


def test_anymal_lstm_actuator_matches_oracle_free_running():
    """Anymal._compute_torques (LSTM actuator net) over 8 free-running substeps, with a reset_idx() of a few
    envs and a fused-step reset in between (both must clear exactly the reset envs' hidden / cell state)."""
    from oracle import legged_oracle as lo
    n = 1501                                  # 18012 actuators: not a multiple of the CTA size
    env = helpers.make_env("anymal_c_flat", n, device=DEV)
    sd, atol = helpers.sea_setup(None, "cpu")
    sea = lo.SeaLstmOracle(n, env.num_actions, sd)
    p = lo.OracleParams.from_env(env, device="cpu")
    frames = synthetic.make_frames(env, 8, seed=21)
    for it, fr in enumerate(frames):
        synthetic.load_frame(env, fr)
        a = fr["actions"].clip(-p.clip_actions, p.clip_actions)
        got = env._compute_torques(a)
        want = sea.compute_torques(a.cpu(), p.action_scale, p.default_dof_pos, env.dof_pos.cpu(), env.dof_vel.cpu())
        assert got.shape == want.shape == (n * env.num_actions,)
        helpers.assert_close_abs(f"sea torques[{it}]", got, want, atol)
        helpers.assert_close(f"sea_input[{it}]", env.sea_input, sea.sea_input)
        helpers.assert_close(f"sea_hidden_state[{it}]", env.sea_hidden_state, sea.sea_hidden_state)
        helpers.assert_close(f"sea_cell_state[{it}]", env.sea_cell_state, sea.sea_cell_state)
        if it == 2:
            ids = torch.tensor([0, 5, 777, n - 1], device=DEV)
            env.reset_idx(ids)
            sea.reset(ids.cpu())
        if it == 4:                            # resets decided inside the fused step kernel
            env.episode_length_buf[10:20] = int(env.max_episode_length)
            env.step(fr["actions"].clone())
            torch.cuda.synchronize()
            k = int(env.step_counts[0])
            assert k >= 10
            sea.reset(env.reset_ids[:k].long().cpu())
            assert float(env.sea_hidden_state_per_env[:, 10:20].abs().sum()) == 0.0
    assert float(env.sea_hidden_state.abs().sum()) > 0.0


def test_python_reward_terms_of_a_subclass_enter_the_fused_kernel():
    """The reference's way of adding a reward: define `_reward_<name>` on a subclass and give it a scale
    (legged_robot.py:588-593).  Such terms run as Python here, on the state lgk_pre_reward_state prepares,
    and are scaled / summed / accumulated by the kernel in their alphabetical slot; a python method named
    like a built-in term replaces it."""
    from legged_gym_b200.envs.go1.go1 import Go1

    class Go1PythonTerms(Go1):
        def _reward_alive_progress(self):      # episode counter, termination flags, commands, body-frame velocity
            return (self.episode_length_buf.float() / self.max_episode_length) * (~self.reset_buf) * \
                torch.exp(-torch.abs(self.commands[:, 0] - self.base_lin_vel[:, 0]))

        def _reward_clearance(self):           # height scan and projected gravity
            return torch.mean(torch.clip(self.root_states[:, 2:3] - 0.3 - self.measured_heights, 0., 1.), dim=1) * \
                (1. + self.projected_gravity[:, 2])

        def _reward_torques(self):             # overrides the built-in term of the same name
            return torch.sum(torch.abs(self.torques), dim=1)

    oracle_terms = {
        "alive_progress": lambda p, s: (s.episode_length_buf.float() / p.max_episode_length) * (~s.reset_buf) *
        torch.exp(-torch.abs(s.commands[:, 0] - s.base_lin_vel[:, 0])),
        "clearance": lambda p, s: torch.mean(torch.clip(s.root_states[:, 2:3] - 0.3 - s.measured_heights, 0., 1.), dim=1) *
        (1. + s.projected_gravity[:, 2]),
        "torques": lambda p, s: torch.sum(torch.abs(s.torques), dim=1),
    }

    def scales(cfg):
        cfg.rewards.scales.alive_progress = 0.7
        cfg.rewards.scales.clearance = -0.3
    n = helpers.run_parity("go1_rough_full", 777, 5, extra=scales, task_class=Go1PythonTerms,
                           oracle_extra_terms=oracle_terms)
    assert n > 0


@pytest.mark.parametrize("sums", [30.0, 15.9992, 15.99995, 3.0])
def test_command_curriculum_widens_the_range_on_the_gated_step(sums):
    """update_command_curriculum (:465-474): on the one step in max_episode_length where the gate is open, good
    velocity tracking of the reset envs widens lin_vel_x by 0.5 BEFORE those envs resample their commands (and
    after the callback's resample, which still uses the old range).  The gate compares the mean episode sum
    INCLUDING this step's reward with 0.8 * scale * max_episode_length = 16: the values next to 16 make the
    decision hinge on that last term."""
    from oracle import legged_oracle as lo

    def curriculum(cfg):
        cfg.commands.curriculum = True
        cfg.commands.max_curriculum = 2.0          # the shipped 1.0 equals the initial range: nothing could widen
    env = helpers.make_env("go1_rough_full", 300, device=DEV, seed=6, extra=curriculum)
    helpers.randomise_env_state(env, 6)
    env.common_step_counter = int(env.max_episode_length) - 1          # this step: counter % max_episode_length == 0
    env.episode_sums["tracking_lin_vel"][:] = sums                     # gate: mean / 1000 > 0.8 * scale (= 0.016)
    env.episode_length_buf[:24] = int(env.max_episode_length)          # 24 time-outs
    frames = synthetic.make_frames(env, 2, seed=41)
    synthetic.load_frame(env, frames[0])
    p = lo.OracleParams.from_env(env, device="cpu")
    s = lo.OracleState.from_env(env, p, device="cpu")
    assert p.command_ranges["lin_vel_x"] == [-1.0, 1.0]
    for it, fr in enumerate(frames):
        taps = synthetic.make_taps(env, seed=600 + it)
        synthetic.load_frame(env, fr)
        env.use_rng_taps(taps, index_math=_lgk.INDEX_MATH_TORCH_CPU)
        env.step(fr["actions"].clone())
        torch.cuda.synchronize()
        s.root_states.copy_(fr["root_states"].cpu()); s.dof_state.copy_(fr["dof_state"].cpu())
        s.contact_forces.copy_(fr["contact_forces"].cpu().view_as(s.contact_forces)); s.torques.copy_(fr["torques"].cpu())
        lo.step(p, s, fr["actions"].cpu().clone(), {k: v.cpu() for k, v in taps.items()})
        helpers.compare_env_to_oracle(env, s, tag=f"[command curriculum step {it}] ")
        got = [float(v) for v in env.command_ranges["lin_vel_x"]]
        assert got == [float(v) for v in p.command_ranges["lin_vel_x"]]
        if sums >= 30.0:
            assert got == [-1.5, 1.5]
        if sums <= 3.0:
            assert got == [-1.0, 1.0]
        assert float(env.extras["episode"]["max_command_x"]) == got[1]


def test_post_physics_step_callback_override_edits_the_commands():
    """A derived task whose callback post-processes the commands (no turning while nearly standing) and keeps a
    gait clock: the override sees this step's resampled commands, the kernel uses what it leaves."""
    from oracle import legged_oracle as lo
    from legged_gym_b200.envs.go1.go1 import Go1

    class Go1Gait(Go1):
        def _init_buffers(self):
            super()._init_buffers()
            self.gait_clock = torch.zeros(self.num_envs, device=self.device)

        def _post_physics_step_callback(self):
            super()._post_physics_step_callback()
            self.commands[:, 2] *= (torch.norm(self.commands[:, :2], dim=1) > 0.5)
            self.gait_clock = (self.gait_clock + self.dt * 2.0) % 1.0

    env = helpers.make_env("go1_rough_full", 512, device=DEV, seed=13, task_class=Go1Gait)
    helpers.randomise_env_state(env, 13)
    frames = synthetic.make_frames(env, 3, seed=73)
    synthetic.load_frame(env, frames[0])
    p = lo.OracleParams.from_env(env, device="cpu")
    s = lo.OracleState.from_env(env, p, device="cpu")
    base_cb = lo.post_physics_step_callback

    def edited(pp, ss, taps):
        base_cb(pp, ss, taps)
        ss.commands[:, 2] *= (torch.norm(ss.commands[:, :2], dim=1) > 0.5)
    lo.post_physics_step_callback = edited
    try:
        for it, fr in enumerate(frames):
            taps = synthetic.make_taps(env, seed=950 + it)
            synthetic.load_frame(env, fr)
            env.use_rng_taps(taps, index_math=_lgk.INDEX_MATH_TORCH_CPU)
            env.step(fr["actions"].clone())
            torch.cuda.synchronize()
            s.root_states.copy_(fr["root_states"].cpu()); s.dof_state.copy_(fr["dof_state"].cpu())
            s.contact_forces.copy_(fr["contact_forces"].cpu().view_as(s.contact_forces)); s.torques.copy_(fr["torques"].cpu())
            lo.step(p, s, fr["actions"].cpu().clone(), {k: v.cpu() for k, v in taps.items()})
            helpers.compare_env_to_oracle(env, s, tag=f"[callback override step {it}] ")
        assert abs(float(env.gait_clock[0]) - (3 * env.dt * 2.0) % 1.0) < 1e-6
        assert int((env.commands[:, 2] == 0).sum()) > 50
    finally:
        lo.post_physics_step_callback = base_cb


def test_check_termination_override_adds_a_condition():
    """A derived task that also terminates on a large tilt (a common customisation of :143-148).  The kernel
    acts on the mask the override leaves -- resets, the termination reward, episode statistics -- which equals
    the oracle run with the same extra condition."""
    from oracle import legged_oracle as lo
    from legged_gym_b200.envs.go1.go1 import Go1

    class Go1TiltTermination(Go1):
        def check_termination(self):
            super().check_termination()
            self.reset_buf = self.reset_buf | (self.projected_gravity[:, 2] > -0.985)      # re-binds, like :146

    env = helpers.make_env("go1_rough_full", 640, device=DEV, seed=12, task_class=Go1TiltTermination)
    helpers.randomise_env_state(env, 12)
    frames = synthetic.make_frames(env, 3, seed=71)
    synthetic.load_frame(env, frames[0])
    p = lo.OracleParams.from_env(env, device="cpu")
    s = lo.OracleState.from_env(env, p, device="cpu")
    base_check = lo.check_termination

    def tilted(pp, ss):
        base_check(pp, ss)
        ss.reset_buf = ss.reset_buf | (ss.projected_gravity[:, 2] > -0.985)
    lo.check_termination = tilted
    try:
        extra_resets = 0
        for it, fr in enumerate(frames):
            taps = synthetic.make_taps(env, seed=900 + it)
            synthetic.load_frame(env, fr)
            env.use_rng_taps(taps, index_math=_lgk.INDEX_MATH_TORCH_CPU)
            addr = env.reset_buf.data_ptr()
            env.step(fr["actions"].clone())
            torch.cuda.synchronize()
            assert env.reset_buf.data_ptr() == addr
            s.root_states.copy_(fr["root_states"].cpu()); s.dof_state.copy_(fr["dof_state"].cpu())
            s.contact_forces.copy_(fr["contact_forces"].cpu().view_as(s.contact_forces)); s.torques.copy_(fr["torques"].cpu())
            lo.step(p, s, fr["actions"].cpu().clone(), {k: v.cpu() for k, v in taps.items()})
            helpers.compare_env_to_oracle(env, s, tag=f"[tilt termination step {it}] ")
            extra_resets += int((s.reset_buf & ~s.time_out_buf).sum())
        assert extra_resets > 20       # the tilt condition fires for a good share of the synthetic orientations
    finally:
        lo.check_termination = base_check


def test_compute_observations_override_appends_columns():
    """The usual way a derived task adds observations: override compute_observations(), call super(), torch.cat.
    The override runs after the fused launch with self.obs_buf holding the kernel's (noisy, clipped) base
    observation; the result is clipped like step() does (:103-104)."""
    from legged_gym_b200.envs.go1.go1 import Go1

    class Go1ExtraObs(Go1):
        def compute_observations(self):
            super().compute_observations()
            contact = (self.contact_forces[:, self.feet_indices, 2] > 1.).float()
            phase = (self.episode_length_buf.float() * 300.).unsqueeze(1)        # exceeds clip_observations
            self.obs_buf = torch.cat((self.obs_buf, contact, phase), dim=-1)

    def wider(cfg):
        cfg.env.num_observations = 235 + 4 + 1
    n = 257
    plain = helpers.make_env("go1_rough_full", n, device=DEV, seed=8)
    extra = helpers.make_env("go1_rough_full", n, device=DEV, seed=8, extra=wider, task_class=Go1ExtraObs)
    assert extra.num_obs == 240 and extra.obs_buf.shape == (n, 240)
    for e in (plain, extra):
        helpers.randomise_env_state(e, 8)
    frames = synthetic.make_frames(plain, 3, seed=55)
    for fr in frames:
        for e in (plain, extra):
            synthetic.load_frame(e, fr)
        o1 = plain.step(fr["actions"].clone())[0]
        o2, _, r2, d2, _ = extra.step(fr["actions"].clone())
        torch.cuda.synchronize()
        assert o2.shape == (n, 240) and extra.get_observations() is o2
        assert torch.equal(o2[:, :235], o1) and torch.equal(r2, plain.rew_buf) and torch.equal(d2, plain.reset_buf)
        assert torch.equal(o2[:, 235:239], (extra.contact_forces[:, extra.feet_indices, 2] > 1.).float())
        want_phase = (extra.episode_length_buf.float() * 300.).clip(max=extra.cfg.normalization.clip_observations)
        assert torch.equal(o2[:, 239], want_phase) and float(o2[:, 239].max()) == 100.0
    with pytest.raises(ValueError, match="num_observations"):
        helpers.make_env("go1_rough_full", 8, device=DEV, extra=wider)      # wider obs without an override


def test_outputs_written_straight_into_rollout_storage():
    """with the env's output pointers moved into [T,N,O] storage the kernel's stores are
    the only pass; the storage must equal what a plain env returns step by step (same seed, in-kernel RNG)."""
    from legged_gym_b200.utils.rollout import RolloutStorageWriter
    n, T = 333, 5
    envs = [helpers.make_env("go1_rough_full", n, device=DEV, seed=9) for _ in range(2)]
    for e in envs:
        helpers.randomise_env_state(e, 9)
    frames = synthetic.make_frames(envs[0], T, seed=77)
    for e in envs:
        synthetic.load_frame(e, frames[0])
    plain, wr = envs[0], RolloutStorageWriter(envs[1], T)
    for rollout in range(2):
        for t in range(T):
            for e in envs:
                synthetic.load_frame(e, frames[t])
            obs, _, rew, reset, _ = plain.step(frames[t]["actions"].clone())
            o2, _, r2, d2, _ = wr.step(frames[t]["actions"].clone())
            torch.cuda.synchronize()
            assert o2.data_ptr() == wr.observations[t + 1].data_ptr()
            assert torch.equal(wr.observations[t + 1], obs) and torch.equal(wr.rewards[t, :, 0], rew)
            assert torch.equal(wr.dones[t, :, 0], reset) and torch.equal(d2, reset)
        with pytest.raises(AssertionError, match="overflow"):
            wr.step(frames[0]["actions"])
        wr.clear()
        assert torch.equal(wr.observations[0], plain.obs_buf)
    with pytest.raises(ValueError):
        plain.bind_output_tensors(obs_buf=torch.zeros(n, 7, device=DEV))


def test_env_state_checkpoint_resumes_bit_identically():
    """A fresh env restored from state_dict() continues exactly like the env that kept running (in-kernel
    Philox RNG is keyed by the step counter, so no generator state has to travel)."""
    n = 700
    a = helpers.make_env("go1_rough_full", n, device=DEV, seed=5)
    helpers.randomise_env_state(a, 5)
    frames = synthetic.make_frames(a, 6, seed=31)
    for t in range(3):
        synthetic.load_frame(a, frames[t])
        a.step(frames[t]["actions"].clone())
    sd = a.state_dict()
    b = helpers.make_env("go1_rough_full", n, device=DEV, seed=5)
    b.load_state_dict(sd)
    for t in range(3, 6):
        for e in (a, b):
            synthetic.load_frame(e, frames[t])
            e.step(frames[t]["actions"].clone())
        torch.cuda.synchronize()
        for name in ("obs_buf", "rew_buf", "reset_buf", "commands", "root_states", "dof_state", "feet_air_time",
                     "episode_length_buf", "terrain_levels", "env_origins", "_episode_sums_buf", "step_counts"):
            assert torch.equal(getattr(a, name), getattr(b, name)), f"step {t}: {name} differs after resume"
        k = int(a.step_counts[0])
        assert torch.equal(a.reset_ids[:k], b.reset_ids[:k])
    assert int(a.reset_buf.sum()) >= 0 and a.common_step_counter == b.common_step_counter


# ------------------------------------------------------------------------------------------- resets
def test_reset_idx_subset_and_reset_all():
    from oracle import legged_oracle as lo
    env = helpers.make_env("go1_rough_full", 500, device=DEV, seed=4)
    helpers.randomise_env_state(env, 4)
    fr = synthetic.make_frames(env, 1, seed=6)[0]
    synthetic.load_frame(env, fr)
    taps = synthetic.make_taps(env, seed=12)
    p = lo.OracleParams.from_env(env, device="cpu")
    s = lo.OracleState.from_env(env, p, device="cpu")
    ids = torch.tensor([0, 7, 31, 32, 33, 250, 499], device=DEV)
    env.use_rng_taps(taps)
    env.reset_idx(ids)
    torch.cuda.synchronize()
    lo.reset_idx(p, s, ids.cpu(), {k: v.cpu() for k, v in taps.items()})
    for name in ("root_states", "dof_state", "commands", "feet_air_time", "env_origins"):
        helpers.assert_close(name, getattr(env, name).reshape(-1), getattr(s, name).reshape(-1))
    helpers.assert_close("episode_length_buf", env.episode_length_buf, s.episode_length_buf, exact=True)
    helpers.assert_close("terrain_levels", env.terrain_levels, s.terrain_levels, exact=True)
    helpers.assert_close("reset_ids", env.reset_ids[:7].long(), ids, exact=True)
    for k, v in s.extras["episode"].items():
        helpers.assert_close(f"extras[{k}]", env.extras["episode"][k].reshape(1), torch.as_tensor(v).float().reshape(1))
    for k in env.episode_sums:
        helpers.assert_close(f"sums[{k}]", env.episode_sums[k], s.episode_sums[k])
    env.reset_idx(torch.zeros(0, dtype=torch.long, device=DEV))       # empty list: early-out (:160)
    obs, priv = env.reset()                                            # reset_idx(all) + zero-action step
    assert obs is env.obs_buf and priv is None and torch.isfinite(obs).all()
    assert (env.episode_length_buf <= 1).all()


# ------------------------------------------------------------------------------------------- tables / probes
def test_height_min3_table():
    env = helpers.make_env("go1_rough_full", 8, device=DEV)
    env._kernels()
    hs = env.height_samples.cpu()
    want = torch.min(torch.min(hs[:-1, :-1], hs[1:, :-1]), hs[:-1, 1:])
    assert torch.equal(env._height_min3.cpu()[:-1, :hs.shape[1] - 1], want)


def test_yaw_quaternion_formula_is_torch_exact():
    lib = _lgk.load()
    g = torch.Generator().manual_seed(0)
    q = torch.randn(200000, 4, generator=g)
    q = (q / q.norm(dim=1, keepdim=True)).to(DEV)
    out = torch.empty(200000, 2, device=DEV)
    _lgk.check("yaw", lib.lgk_debug_yaw_quat(_lgk.ptr(q), 200000, 1, _lgk.ptr(out), _lgk.current_stream(DEV)))
    yq = q.clone()
    yq[:, :2] = 0
    for dev in ("cpu", DEV):
        y = yq.to(dev)
        want = (y / y.norm(p=2, dim=-1).clamp(min=1e-9).unsqueeze(-1))[:, 2:]
        assert torch.equal(out.to(dev), want), f"normalize([0,0,qz,qw]) differs from torch on {dev}"


# ------------------------------------------------------------------------------------------- (3) full size
def _full_env(seed, n=65536):
    env = helpers.make_env("go1_rough_full", n, device=DEV, seed=seed)
    g = torch.Generator().manual_seed(77)
    env.episode_length_buf = torch.randint(0, 1002, (n,), generator=g).to(DEV)
    return env


@pytest.mark.parametrize("N", [65536, 65552])   # 65 552: a multiple of 16 but not of 64 (the tail kernel's last unit is ragged)
def test_full_size_properties_with_in_kernel_rng(N):
    from oracle import legged_oracle as lo
    env = _full_env(1, N)
    frames = synthetic.make_frames(env, 2, seed=21)
    p = lo.OracleParams.from_env(env, device=DEV)
    n_resets = 0
    for it in range(4):
        fr = frames[it % 2]
        synthetic.load_frame(env, fr)
        ep_before = env.episode_length_buf.clone()
        s = lo.OracleState.from_env(env, p, device=DEV)
        env.step(fr["actions"].clone())
        torch.cuda.synchronize()
        # terminations / id lists: exact against the oracle's check_termination on the same inputs
        s.episode_length_buf += 1
        lo.check_termination(p, s)
        assert torch.equal(env.reset_buf, s.reset_buf) and torch.equal(env.time_out_buf, s.time_out_buf)
        nr, ns = int(env.step_counts[0]), int(env.step_counts[1])
        ids = env.reset_buf.nonzero().flatten()
        assert nr == len(ids) and torch.equal(env.reset_ids[:nr].long(), ids)          # ascending, complete
        rs = ((ep_before + 1) % 500 == 0).nonzero().flatten()
        assert ns == len(rs) and torch.equal(env.resample_ids[:ns].long(), rs)
        # height-field indices at full size: measured_heights == int16 table * 0.005f exactly
        s.root_states.copy_(fr["root_states"])
        want_h = lo.get_heights(p, s)
        assert torch.equal(env.measured_heights, want_h)
        # episode bookkeeping
        assert (env.episode_length_buf[ids] == 0).all()
        keep = ~env.reset_buf
        assert torch.equal(env.episode_length_buf[keep], ep_before[keep] + 1)
        assert (env.feet_air_time[ids] == 0).all() and (env.dof_vel[ids] == 0).all()
        for k in env.episode_sums:
            assert (env.episode_sums[k][ids] == 0).all()
        # reset poses: within 1 m of the (new) platform centre, joints in default * [0.5, 1.5]
        d = (env.root_states[ids, :2] - env.env_origins[ids, :2]).abs()
        assert (d <= 1.0 + 1e-5).all() and (env.root_states[ids, 7:13].abs() <= 0.5).all()
        ratio = env.dof_pos[ids] / env.default_dof_pos
        assert (ratio >= 0.5 - 1e-5).all() and (ratio <= 1.5 + 1e-5).all()
        # observations: finite, clipped, noise bounded by the noise vector around the noiseless value
        assert torch.isfinite(env.obs_buf).all() and (env.obs_buf.abs() <= 100).all()
        clean = torch.clip(env.root_states[:, 2:3] - 0.5 - env.measured_heights, -1, 1) * 5.0
        dev_h = (env.obs_buf[:, 48:] - clean).abs()
        assert (dev_h[keep] <= 0.5 + 1e-5).all() and dev_h[keep].mean() > 0.2        # U(-0.5, 0.5): mean |x| = 0.25
        noise_h = (env.obs_buf[:, 48:] - clean)[keep]
        assert abs(float(noise_h.mean())) < 2e-3 and abs(float(noise_h.var()) - 1 / 12) < 2e-3
        assert torch.equal(env.obs_buf[:, 36:48], env.actions)                        # zero noise scale columns
        assert torch.equal(env.last_actions, env.actions) and torch.equal(env.last_dof_vel, env.dof_vel)
        assert (env.rew_buf[keep] >= 0).all()                                         # only_positive_rewards
        # every env's sums were visited, the ragged last unit included: exp(-err / sigma) * scale * dt > 0 each step
        assert torch.isfinite(env.rew_buf).all()
        assert (env.episode_sums["tracking_lin_vel"][keep] > 0).all() and (env.episode_sums["torques"][keep] < 0).all()
        n_resets += nr
    assert n_resets > 1000


def test_in_kernel_rng_is_deterministic_per_seed_and_decorrelated_across_steps():
    outs = []
    for seed in (1, 1, 2):
        env = _full_env(seed, n=8192)
        frames = synthetic.make_frames(env, 1, seed=21)
        rows = []
        for it in range(2):
            synthetic.load_frame(env, frames[0])
            env.episode_length_buf = torch.full((8192,), 10 + it, device=DEV)
            env.step(frames[0]["actions"].clone())
            rows.append(env.obs_buf.clone())
        outs.append(rows)
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    assert not torch.equal(outs[0][0][:, :3], outs[2][0][:, :3])
    a = outs[0][0][:, 48:] - outs[0][1][:, 48:]            # same state, different step -> different noise
    assert (a != 0).float().mean() > 0.99
    n0 = (outs[0][0][:, 100] - outs[2][0][:, 100])
    n1 = (outs[0][1][:, 100] - outs[2][1][:, 100])
    assert abs(float(torch.corrcoef(torch.stack([n0, n1]))[0, 1])) < 0.05


def test_step_is_graph_capturable_and_sync_free():
    env = helpers.make_env("go1_rough_full", 4096, device=DEV, seed=9)
    frames = synthetic.make_frames(env, 1, seed=2)
    synthetic.load_frame(env, frames[0])
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for _ in range(3):
            env.step(frames[0]["actions"])
        st.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            env.step(frames[0]["actions"])
        ep0 = env.episode_length_buf.clone()
        for _ in range(5):
            g.replay()
        st.synchronize()
    keep = env.episode_length_buf > 4
    assert keep.any() and torch.equal(env.episode_length_buf[keep], ep0[keep] + 5)


def test_actuator_net_is_graph_capturable():
    """The tcgen05 MLP's per-device workspace is allocated by the env's set-up call (lgk_actuator_advance with zero
    envs), so the very first substep may already sit inside a CUDA-graph capture; replays equal eager calls bit
    for bit (history shift included)."""
    n = 777
    eager = helpers.make_env("go1_rough_actuator", n, device=DEV)
    graphed = helpers.make_env("go1_rough_actuator", n, device=DEV)
    fr = synthetic.make_frames(eager, 1, seed=5)[0]
    synthetic.load_frame(eager, fr)
    synthetic.load_frame(graphed, fr)
    a = fr["actions"].clone()
    stream = torch.cuda.Stream(DEV)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(stream):
        with torch.cuda.graph(g, stream=stream):
            graphed.actuator_advance(a)          # first call with envs on this env: inside the capture
    torch.cuda.synchronize()
    assert float(graphed.dVel.abs().sum()) == 0.0   # capture records, it does not run
    for it in range(3):
        want = eager.actuator_advance(a).clone()
        g.replay()
        torch.cuda.synchronize()
        assert torch.equal(graphed.dVel, want), it
        assert torch.equal(graphed.pos_err_buffs, eager.pos_err_buffs)


def test_actuator_two_weight_sets_on_two_streams():
    """Two envs with different actuator-net weights on TWO streams of one device, many substeps in flight at once:
    each env owns its weight-image workspace and every launch carries its biases / output layer in its own
    parameter block, so nothing is shared (round 1 kept one image buffer and one __constant__ block per device)."""
    from legged_gym_b200.envs.go1.go1 import random_mlp_state_dict
    from oracle import legged_oracle as lo
    n = 8192

    def with_seed(seed):
        def edit(cfg):
            cfg.control.actuator_net_random_seed = seed
        return edit
    envs = [helpers.make_env("go1_rough_actuator", n, device=DEV, extra=with_seed(s)) for s in (0, 7)]
    oracles = [lo.ActuatorOracle(n, random_mlp_state_dict(s)) for s in (0, 7)]
    atols = [helpers.actuator_dvel_atol(random_mlp_state_dict(s)) for s in (0, 7)]
    assert envs[0]._actuator_workspace.data_ptr() != envs[1]._actuator_workspace.data_ptr()
    fr = synthetic.make_frames(envs[0], 1, seed=12)[0]
    for env in envs:
        synthetic.load_frame(env, fr)
    streams = [torch.cuda.Stream(DEV), torch.cuda.Stream(DEV)]
    torch.cuda.synchronize()
    reps = 6
    for k in range(reps):                      # interleaved issue: the two streams' kernels overlap on the GPU
        for env, st in zip(envs, streams):
            with torch.cuda.stream(st):
                env.actuator_advance(fr["actions"])
    torch.cuda.synchronize()
    for env, act, atol in zip(envs, oracles, atols):
        for k in range(reps):
            want = act.advance(fr["actions"].cpu(), env.dof_pos.cpu(), env.dof_vel.cpu())
        helpers.assert_close_abs("dVel after 6 overlapped substeps", env.dVel, want, atol)


def test_actuator_two_weight_sets_alternate():
    """Two envs with different actuator-net weights take turns on one stream: the weight images of the tcgen05 kernel
    are rebuilt on every call, behind a programmatic launch dependency, and must never be stale."""
    from legged_gym_b200.envs.go1.go1 import random_mlp_state_dict
    from oracle import legged_oracle as lo
    n = 640

    def with_seed(seed):
        def edit(cfg):
            cfg.control.actuator_net_random_seed = seed
        return edit
    envs = [helpers.make_env("go1_rough_actuator", n, device=DEV, extra=with_seed(s)) for s in (0, 7)]
    oracles = [lo.ActuatorOracle(n, random_mlp_state_dict(s)) for s in (0, 7)]
    atols = [helpers.actuator_dvel_atol(random_mlp_state_dict(s)) for s in (0, 7)]
    assert not torch.equal(envs[0].actuator_weights, envs[1].actuator_weights)
    frames = synthetic.make_frames(envs[0], 4, seed=12)
    for it, fr in enumerate(frames):
        for env in envs:
            synthetic.load_frame(env, fr)
        got = [env.actuator_advance(fr["actions"]) for env in envs]      # back to back, no sync in between
        for env, act, atol, g in zip(envs, oracles, atols, got):
            want = act.advance(fr["actions"].cpu(), env.dof_pos.cpu(), env.dof_vel.cpu())
            helpers.assert_close_abs(f"dVel[{it}]", g, want, atol)
