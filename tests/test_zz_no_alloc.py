"""lgk_actuator_advance_static: the allocation-free variant of the Go1 actuator substep (scratch in a file-scope
device array inside liblgk.so).  It has no counterpart in the reference; the check is against the existing entry
point lgk_actuator_advance on the same inputs, through the C-ABI, on two calls in a row (the second call meets the
first call's contents in the shared scratch array and a shifted history).

Collected last on purpose; it passes alone in a fresh process and after the
other tests."""
import ctypes as C

import pytest
import torch

import helpers
from legged_gym_b200 import _lgk

DEV = "cuda:0"


def _inputs(n, seed):
    from legged_gym_b200.envs.go1 import go1
    g = torch.Generator().manual_seed(seed)
    blob = go1.pack_mlp_weights(go1.random_mlp_state_dict(0))     # W0[128,30] b0 W1 b1 W2 b2 W3[3,128] b3
    assert blob.numel() == 128 * 30 + 128 + 2 * (128 * 128 + 128) + 3 * 128 + 3
    norm = torch.cat([torch.tensor(v).tile((go1.LEG_NUM,))
                      for v in (go1.POS_ERR_MEAN, go1.POS_ERR_STD, go1.VEL_MEAN, go1.VEL_STD)]).float()
    calls = [(torch.randn(n, 12, generator=g).float(), (torch.randn(n, 12, 2, generator=g) * 0.7).float())
             for _ in range(2)]
    return blob, norm, calls


def _run(entry, n, blob, norm, calls, kernel_env, monkeypatch):
    """both entry points on private copies of the state: returns dVel of each call and the final histories"""
    lib = _lgk.load()
    monkeypatch.delenv("LGK_MLP_FFMA", raising=False)
    if kernel_env is None:
        monkeypatch.delenv("LGK_MLP_KERNEL", raising=False)
    else:
        monkeypatch.setenv("LGK_MLP_KERNEL", kernel_env)
    P = _lgk.ptr
    w, nm = blob.to(DEV), norm.to(DEV)
    pe = torch.zeros(n, 12, 5, device=DEV)
    vh = torch.zeros(n, 12, 5, device=DEV)
    mi = torch.full((n, 120), float("nan"), device=DEV)
    ws = torch.zeros((int(lib.lgk_actuator_workspace_bytes()) + 7) // 8, dtype=torch.int64, device=DEV)
    st = _lgk.current_stream(DEV)
    outs = []
    for actions, dof_state in calls:
        a, ds = actions.to(DEV), dof_state.to(DEV)
        dvel = torch.full((n, 12), float("nan"), device=DEV)
        if entry == "static":
            rc = lib.lgk_actuator_advance_static(n, P(a), P(ds), P(pe), P(vh), P(w), P(nm), P(dvel), st)
        else:
            rc = lib.lgk_actuator_advance(n, P(a), P(ds), P(pe), P(vh), P(mi), P(w), P(nm), P(dvel), P(ws), st)
        _lgk.check(entry, rc)
        torch.cuda.synchronize()
        outs.append(dvel.cpu())
    return outs, pe.cpu(), vh.cpu()


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1000, _lgk.LGK_STATIC_MAX_ENVS])
def test_static_entry_matches_allocating_entry_twice_in_a_row(n, monkeypatch):
    from legged_gym_b200.envs.go1.go1 import random_mlp_state_dict
    blob, norm, calls = _inputs(n, seed=3)
    got, pe_s, vh_s = _run("static", n, blob, norm, calls, None, monkeypatch)
    # same arithmetic as the existing entry point's fp32 FFMA kernel: bit-equal
    ffma, pe_f, vh_f = _run("lgk_actuator_advance", n, blob, norm, calls, "ffma", monkeypatch)
    # and within the suite's actuator tolerance of the existing entry point's default (tcgen05) kernel
    dflt, pe_d, vh_d = _run("lgk_actuator_advance", n, blob, norm, calls, None, monkeypatch)
    atol = helpers.actuator_dvel_atol(random_mlp_state_dict(0))
    for it in range(2):
        assert torch.isfinite(got[it]).all()
        assert torch.equal(got[it], ffma[it]), f"call {it}: static entry differs from the FFMA kernel of lgk_actuator_advance"
        helpers.assert_close_abs(f"dVel[{it}] vs default kernel", got[it], dflt[it], atol)
    assert torch.equal(pe_s, pe_f) and torch.equal(vh_s, vh_f)
    assert torch.equal(pe_s, pe_d) and torch.equal(vh_s, vh_d)
    assert not torch.equal(got[0], got[1])


@pytest.mark.gpu
def test_static_entry_allocates_nothing_and_rejects_oversize():
    lib = _lgk.load()
    n = 512
    blob, norm, calls = _inputs(n, seed=5)
    P = _lgk.ptr
    w, nm = blob.to(DEV), norm.to(DEV)
    pe, vh = torch.zeros(n, 12, 5, device=DEV), torch.zeros(n, 12, 5, device=DEV)
    a, ds = calls[0][0].to(DEV), calls[0][1].to(DEV)
    dvel = torch.empty(n, 12, device=DEV)
    st = _lgk.current_stream(DEV)
    assert lib.lgk_actuator_advance_static(0, P(a), P(ds), P(pe), P(vh), P(w), P(nm), P(dvel), st) == 0   # set-up call
    # one warm call first: the driver loads the two kernels' code lazily on their first launch
    assert lib.lgk_actuator_advance_static(n, P(a), P(ds), P(pe), P(vh), P(w), P(nm), P(dvel), st) == 0
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        assert lib.lgk_actuator_advance_static(n, P(a), P(ds), P(pe), P(vh), P(w), P(nm), P(dvel), st) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, "the allocation-free entry point changed the device's free memory"
    too_many = _lgk.LGK_STATIC_MAX_ENVS + 1
    rc = lib.lgk_actuator_advance_static(too_many, P(a), P(ds), P(pe), P(vh), P(w), P(nm), P(dvel), st)
    assert rc == -8
    with pytest.raises(_lgk.LgkError, match="LGK_ERR_CAPACITY"):
        _lgk.check("lgk_actuator_advance_static", rc)


def test_static_entry_argument_checks_need_no_gpu():
    lib = _lgk.load()
    one, null = C.c_void_p(16), C.c_void_p(None)
    assert lib.lgk_actuator_advance_static(_lgk.LGK_STATIC_MAX_ENVS + 1, one, one, one, one, one, one, one, null) == -8
    assert lib.lgk_actuator_advance_static(-1, one, one, one, one, one, one, one, null) == -2
    assert lib.lgk_actuator_advance_static(4, one, null, one, one, one, one, one, null) == -1
    assert lib.lgk_actuator_advance_static(4, one, C.c_void_p(20), one, one, one, one, one, null) == -5
