"""ctypes binding of liblgk.so (the C ABI in ``include/lgk.h``).

There is no fallback: if the shared library has not been built the import of the kernels fails
loudly (``LgkLibraryMissing``), and every entry point raises on a non-zero status.
"""
from __future__ import annotations

import ctypes as C
import os

LGK_ABI_VERSION = 6
LGK_MAX_DOF = 16
LGK_MAX_FEET = 8
LGK_MAX_PEN = 16
LGK_MAX_TERM = 8
LGK_MAX_REWARDS = 32
LGK_MAX_HIP = 4

# LgkControlType
CTRL = {"P": 0, "V": 1, "T": 2}
# LgkRewardTerm -- alphabetical, one per _reward_<name> of the reference
REWARD_TERMS = [
    "action_rate", "ang_vel_xy", "base_height", "collision", "dof_acc", "dof_pos_limits", "dof_vel",
    "dof_vel_limits", "energy", "feet_air_time", "feet_contact_forces", "hip_motion", "lin_vel_z", "no_fly",
    "orientation", "stand_still", "stumble", "torque_limits", "torques", "tracking_ang_vel", "tracking_lin_vel",
]
REWARD_ID = {n: i for i, n in enumerate(REWARD_TERMS)}
REW_EXTERNAL = len(REWARD_TERMS)

INDEX_DIV, INDEX_MUL_RCP = 0, 1
# (p / horizontal_scale).long(): torch-CPU divides, torch-CUDA multiplies by the fp32 reciprocal
# (measured on B200: tools/calibrate_index_math.py)
INDEX_MATH_TORCH_CPU = INDEX_DIV
INDEX_MATH_TORCH_CUDA = INDEX_MUL_RCP

RNG_PHILOX, RNG_TAPS = 0, 1

F_HEADING_COMMAND = 1 << 0
F_MEASURE_HEIGHTS = 1 << 1
F_TERRAIN_PLANE = 1 << 2
F_ADD_NOISE = 1 << 3
F_PUSH_ROBOTS = 1 << 4
F_TERRAIN_CURRICULUM = 1 << 5
F_CUSTOM_ORIGINS = 1 << 6
F_ONLY_POSITIVE = 1 << 7
F_HAS_TERMINATION = 1 << 8
F_INIT_DONE = 1 << 9
F_EP_INCREMENTED = 1 << 10
F_RESET_GIVEN = 1 << 11
F_CALLBACK_GIVEN = 1 << 12
F_FILTERED_CONTACTS = 1 << 13

STATUS = {0: "LGK_OK", -1: "LGK_ERR_NULL", -2: "LGK_ERR_DIM", -3: "LGK_ERR_CONTROL_TYPE", -4: "LGK_ERR_REWARD",
          -5: "LGK_ERR_ALIGN", -6: "LGK_ERR_VERSION", -7: "LGK_ERR_TERRAIN",
          -8: "LGK_ERR_CAPACITY"}


class LgkRewardEntry(C.Structure):
    _fields_ = [("term", C.c_int32), ("scale", C.c_float)]


_F = C.c_float
_I = C.c_int32


class LgkConsts(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("abi_version", C.c_uint32),
        ("num_envs", _I), ("num_dof", _I), ("num_bodies", _I), ("num_feet", _I), ("num_pen", _I), ("num_term", _I),
        ("num_height_points", _I), ("num_obs", _I), ("num_rewards", _I), ("num_commands", _I),
        ("env_id_offset", _I), ("flags", C.c_uint32), ("control_type", _I), ("index_math", _I), ("rng_mode", _I),
        ("envs_per_cta", _I),
        ("dt", _F), ("sim_dt", _F),
        ("max_episode_length", C.c_int64), ("resample_interval", C.c_int64), ("push_interval", C.c_int64),
        ("max_episode_length_s", _F), ("max_push_vel_xy", _F),
        ("clip_actions", _F), ("clip_obs", _F),
        ("obs_scale_lin_vel", _F), ("obs_scale_ang_vel", _F), ("obs_scale_dof_pos", _F), ("obs_scale_dof_vel", _F),
        ("obs_scale_height", _F),
        ("commands_scale", _F * 3),
        ("action_scale", _F),
        ("p_gains", _F * LGK_MAX_DOF), ("d_gains", _F * LGK_MAX_DOF), ("torque_limits", _F * LGK_MAX_DOF),
        ("default_dof_pos", _F * LGK_MAX_DOF), ("dof_pos_lo", _F * LGK_MAX_DOF), ("dof_pos_hi", _F * LGK_MAX_DOF),
        ("dof_vel_limits", _F * LGK_MAX_DOF),
        ("soft_dof_vel_limit", _F), ("soft_torque_limit", _F), ("base_height_target", _F), ("tracking_sigma", _F),
        ("max_contact_force", _F), ("termination_scale", _F),
        ("rewards", LgkRewardEntry * LGK_MAX_REWARDS),
        ("feet_idx", _I * LGK_MAX_FEET), ("pen_idx", _I * LGK_MAX_PEN), ("term_idx", _I * LGK_MAX_TERM),
        ("hip_idx", _I * LGK_MAX_HIP), ("num_hip", _I),
        ("base_init_state", _F * 13),
        ("border_size", _F), ("horizontal_scale", _F), ("vertical_scale", _F), ("half_env_length", _F),
        ("terrain_rows", _I), ("terrain_cols", _I), ("terrain_pitch", _I), ("max_terrain_level", _I),
        ("num_terrain_types", _I),
        ("seed", C.c_uint64),
    ]


_P = C.c_void_p
BUFFER_FIELDS = [
    "root_states", "dof_state", "contact_forces", "torques",
    "actions", "last_actions", "last_dof_vel", "last_root_vel", "commands", "feet_air_time", "episode_sums",
    "episode_length_buf", "terrain_levels", "terrain_types", "env_origins",
    "obs_buf", "rew_buf", "reset_buf", "time_out_buf", "base_lin_vel", "base_ang_vel", "projected_gravity",
    "measured_heights", "reset_ids", "resample_ids", "counts", "episode_stats", "extras", "ext_rewards",
    "height_min3", "height_points_xy", "noise_scale_vec", "terrain_origins", "command_ranges",
    "tap_cmd_callback", "tap_cmd_reset", "tap_push", "tap_dof", "tap_root_xy", "tap_root_vel", "tap_noise",
    "tap_level", "last_contacts", "workspace",
]


class LgkBuffers(C.Structure):
    _fields_ = [(name, _P) for name in BUFFER_FIELDS]


LGK_LSTM_HID = 8
LGK_STATIC_MAX_ENVS = 4096   # lgk_actuator_advance_static: envs its in-library scratch array holds


class LgkLstmWeights(C.Structure):
    """torch.nn.LSTM(2, 8, num_layers=2) + Linear(8, 1) + in/out scales (include/lgk.h)."""
    _G = 4 * LGK_LSTM_HID
    _fields_ = [("w_ih0", _F * (2 * _G)), ("w_hh0", _F * (LGK_LSTM_HID * _G)), ("b_ih0", _F * _G), ("b_hh0", _F * _G),
                ("w_ih1", _F * (LGK_LSTM_HID * _G)), ("w_hh1", _F * (LGK_LSTM_HID * _G)), ("b_ih1", _F * _G),
                ("b_hh1", _F * _G), ("lin_w", _F * LGK_LSTM_HID), ("lin_b", _F), ("in_scale", _F * 2), ("out_scale", _F)]


class LgkSubTerrain(C.Structure):
    _fields_ = [("type", _I), ("row", _I), ("col", _I), ("p", _I * 10), ("draws_off", _I)]


class LgkTerrainGrid(C.Structure):
    _fields_ = [("num_rows", _I), ("num_cols", _I), ("sub_len", _I), ("sub_wid", _I), ("gen_w", _I), ("gen_l", _I),
                ("border", _I), ("tot_rows", _I), ("tot_cols", _I), ("patch_x1", _I), ("patch_x2", _I),
                ("patch_y1", _I), ("patch_y2", _I), ("vertical_scale", C.c_double)]


TER_FLAT, TER_SLOPE, TER_SLOPE_NOISE, TER_STAIRS, TER_OBSTACLES, TER_STONES, TER_GAP, TER_PIT = range(8)


class LgkLibraryMissing(ImportError):
    pass


class LgkError(RuntimeError):
    def __init__(self, fn, code):
        self.code = code
        what = STATUS.get(code, f"cudaError {code}" if code > 0 else f"status {code}")
        super().__init__(f"{fn} failed: {what}")


_LIB = None
LIB_PATH = os.environ.get("LGK_LIB_PATH") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "liblgk.so")


def lib_path() -> str:
    return LIB_PATH


def load():
    """Load liblgk.so once; raise if it is not built (no CPU / torch fallback exists)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.isfile(LIB_PATH):
        raise LgkLibraryMissing(
            f"{LIB_PATH} is not built. Run `python -c 'import __graft_entry__ as g; g.build()'` or "
            f"legged_gym_b200/csrc/build.sh (nvcc, sm_100a). There is no fallback path.")
    lib = C.CDLL(LIB_PATH)
    cp, bp, fp = C.POINTER(LgkConsts), C.POINTER(LgkBuffers), C.c_void_p
    sig = {
        "lgk_abi_version": (C.c_int32, []),
        "lgk_sizeof_consts": (C.c_int64, []),
        "lgk_sizeof_buffers": (C.c_int64, []),
        "lgk_launch_count": (C.c_int64, []),
        "lgk_workspace_bytes": (C.c_int64, [C.c_int32, C.c_int32]),
        "lgk_build_height_min3": (C.c_int, [fp, C.c_int32, C.c_int32, fp, C.c_int32, fp]),
        "lgk_generate_terrain": (C.c_int, [C.POINTER(LgkTerrainGrid), fp, C.c_int32, fp, fp, fp, fp, fp]),
        "lgk_clip_actions": (C.c_int, [fp, fp, C.c_int64, C.c_float, fp]),
        "lgk_compute_poses": (C.c_int, [cp, fp, fp, fp]),
        "lgk_clip_actions_poses": (C.c_int, [cp, fp, fp, fp, fp]),
        "lgk_compute_torques": (C.c_int, [cp, fp, fp, fp, fp, fp]),
        "lgk_actuator_workspace_bytes": (C.c_int64, []),
        "lgk_actuator_advance": (C.c_int, [C.c_int32, fp, fp, fp, fp, fp, fp, fp, fp, fp, fp]),
        "lgk_actuator_advance_static": (C.c_int, [C.c_int32, fp, fp, fp, fp, fp, fp, fp, fp]),
        "lgk_lstm_actuator_advance": (C.c_int, [cp, C.POINTER(LgkLstmWeights), fp, fp, fp, fp, fp, fp, fp]),
        "lgk_lstm_reset_hidden": (C.c_int, [C.c_int32, C.c_int32, fp, fp, C.c_int32, fp, fp, fp]),
        "lgk_post_physics_step": (C.c_int, [cp, bp, fp]),
        "lgk_pre_reward_state": (C.c_int, [cp, bp, fp]),
        "lgk_reset_idx": (C.c_int, [cp, bp, fp, C.c_int32, fp]),
        "lgk_reset_all": (C.c_int, [cp, bp, fp]),
        "lgk_set_step_counter": (C.c_int, [fp, C.c_int64, fp]),
        "lgk_debug_yaw_quat": (C.c_int, [fp, C.c_int32, C.c_int32, fp, fp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)   # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.lgk_abi_version() != LGK_ABI_VERSION:
        raise LgkLibraryMissing(f"{LIB_PATH}: ABI {lib.lgk_abi_version()} != binding {LGK_ABI_VERSION}; rebuild")
    if lib.lgk_sizeof_consts() != C.sizeof(LgkConsts) or lib.lgk_sizeof_buffers() != C.sizeof(LgkBuffers):
        raise LgkLibraryMissing(
            f"struct layout mismatch: C {lib.lgk_sizeof_consts()}/{lib.lgk_sizeof_buffers()} vs "
            f"ctypes {C.sizeof(LgkConsts)}/{C.sizeof(LgkBuffers)}; rebuild or fix _lgk.py")
    _LIB = lib
    return lib


EXPORTED_SYMBOLS = [
    "lgk_abi_version", "lgk_sizeof_consts", "lgk_sizeof_buffers", "lgk_launch_count", "lgk_workspace_bytes",
    "lgk_build_height_min3", "lgk_generate_terrain", "lgk_clip_actions", "lgk_compute_poses", "lgk_clip_actions_poses", "lgk_compute_torques", "lgk_actuator_advance", "lgk_actuator_workspace_bytes",
    "lgk_actuator_advance_static",
    "lgk_lstm_actuator_advance", "lgk_lstm_reset_hidden", "lgk_post_physics_step", "lgk_pre_reward_state", "lgk_reset_idx", "lgk_reset_all", "lgk_set_step_counter", "lgk_debug_yaw_quat",
]


def check(fn_name: str, code: int):
    if code != 0:
        raise LgkError(fn_name, code)


def ptr(t):
    """Raw device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def current_stream(device=None):
    import torch
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)
