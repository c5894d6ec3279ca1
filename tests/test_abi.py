"""CPU tier: the C-ABI library loads, exports every symbol include/lgk.h declares, agrees with the ctypes
mirror on struct layout, and rejects bad arguments with the documented status codes (argument checks run
before any CUDA call, so no GPU is needed)."""
import ctypes as C
import os
import re

import pytest

from legged_gym_b200 import _lgk

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "lgk.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(lgk_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(_lgk.lib_path())
    names = declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/lgk.h but not exported by liblgk.so"
    assert set(names) == set(_lgk.EXPORTED_SYMBOLS), "ctypes binding and header disagree"


def test_header_citations_point_at_the_reference():
    src = open(HEADER).read()
    assert "legged_robot.py:" in src and "go1.py:" in src


def test_struct_layout_and_version():
    lib = _lgk.load()
    assert lib.lgk_abi_version() == _lgk.LGK_ABI_VERSION
    assert lib.lgk_sizeof_consts() == C.sizeof(_lgk.LgkConsts)
    assert lib.lgk_sizeof_buffers() == C.sizeof(_lgk.LgkBuffers)
    hdr = open(HEADER).read()
    assert int(re.search(r"#define LGK_ABI_VERSION (\d+)", hdr).group(1)) == _lgk.LGK_ABI_VERSION
    for name, val in (("LGK_MAX_DOF", _lgk.LGK_MAX_DOF), ("LGK_MAX_FEET", _lgk.LGK_MAX_FEET),
                      ("LGK_MAX_REWARDS", _lgk.LGK_MAX_REWARDS)):
        assert int(re.search(rf"#define {name}\s+(\d+)", hdr).group(1)) == val
    # reward ids follow the header's enum order
    enum = re.search(r"typedef enum LgkRewardTerm \{(.*?)\}", hdr, flags=re.S).group(1)
    terms = [t.lower()[len("lgk_rew_"):] for t in re.findall(r"(LGK_REW_[A-Z_]+)", enum)]
    assert terms[:len(_lgk.REWARD_TERMS)] == _lgk.REWARD_TERMS and terms[len(_lgk.REWARD_TERMS)] == "external"


def _consts(**kw):
    c = _lgk.LgkConsts()
    c.struct_size = C.sizeof(_lgk.LgkConsts)
    c.abi_version = _lgk.LGK_ABI_VERSION
    c.num_envs, c.num_dof = 8, 12
    for k, v in kw.items():
        setattr(c, k, v)
    return c


def test_argument_errors_are_status_codes_not_crashes():
    lib = _lgk.load()
    null = C.c_void_p(None)
    one = C.c_void_p(16)     # never dereferenced: every call below fails validation first
    assert lib.lgk_clip_actions(null, null, 4, 1.0, null) == -1                       # LGK_ERR_NULL
    assert lib.lgk_clip_actions(one, one, -1, 1.0, null) == -2                        # LGK_ERR_DIM
    assert lib.lgk_compute_poses(None, one, one, null) == -1
    assert lib.lgk_compute_poses(C.byref(_consts(abi_version=1)), one, one, null) == -6   # LGK_ERR_VERSION
    assert lib.lgk_compute_poses(C.byref(_consts(num_dof=99)), one, one, null) == -2
    assert lib.lgk_compute_torques(C.byref(_consts(control_type=7)), one, one, one, one, null) == -3   # NameError (:391)
    assert lib.lgk_compute_torques(C.byref(_consts(control_type=1)), one, one, null, one, null) == -1  # V needs last_dof_vel
    assert lib.lgk_post_physics_step(C.byref(_consts()), None, null) == -1
    b = _lgk.LgkBuffers()
    assert lib.lgk_post_physics_step(C.byref(_consts(num_bodies=17, num_commands=4, num_obs=48)), C.byref(b), null) == -2
    assert lib.lgk_actuator_advance(-1, one, one, one, one, one, one, one, one, null, null) == -2   # 0 envs = set-up call
    assert lib.lgk_actuator_advance(4, one, one, one, one, one, one, one, one, C.c_void_p(64), null) == -5   # workspace alignment
    assert lib.lgk_actuator_workspace_bytes() >= 2 * 128 * 32 * 4 + 4 * 128 * 128 * 4
    w = _lgk.LgkLstmWeights()
    assert lib.lgk_lstm_actuator_advance(C.byref(_consts()), None, one, one, null, one, one, one, null) == -1
    assert lib.lgk_lstm_actuator_advance(C.byref(_consts()), C.byref(w), one, one, null, C.c_void_p(20), one, one, null) == -5
    assert lib.lgk_lstm_reset_hidden(8, 12, null, null, 9, one, one, null) == -2      # more ids than envs
    assert lib.lgk_lstm_reset_hidden(8, 12, null, null, 0, one, one, null) == 0       # empty list: no launch (anymal.py:56)
    assert C.sizeof(w) == 4 * (64 + 256 + 64 + 256 + 256 + 64 + 8 + 1 + 2 + 1)
    assert lib.lgk_build_height_min3(null, 4, 4, null, 8, null) == -1
    assert lib.lgk_build_height_min3(one, 1, 4, one, 8, null) == -2
    assert lib.lgk_workspace_bytes(65536, 21) >= 512 + 2 * 4 * (65536 // 32)
    with pytest.raises(_lgk.LgkError, match="LGK_ERR_CONTROL_TYPE"):
        _lgk.check("lgk_compute_torques", -3)


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lgk, "_LIB", None)
    monkeypatch.setattr(_lgk, "LIB_PATH", "/nonexistent/liblgk.so")
    with pytest.raises(_lgk.LgkLibraryMissing, match="no fallback"):
        _lgk.load()
