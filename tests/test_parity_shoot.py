"""Parity of multiple-shooting defects and Jacobian blocks against the oracle (dual numbers in C, themselves
cross-checked against complex-step differentiation of the pinned numpy forward map in test_oracle_golden.py)."""
import numpy as np
import pytest
import torch

from cocofest_b200 import fes
from oracle import fes_oracle as fo
from oracle import fes_oracle_c as foc
from tests.conftest import BACKENDS
from tests.problems import MODELS, P_COLS, PF_COLS, entry_err, oracle_win, random_problem, rel_err, to_lib

RTOL64 = 1e-9
RTOL32 = 1e-4
JKEEP = (2, 3)  # Jacobian blocks [B, ns, row, column]: every (row, column) entry on its own scale


def _X(pb, seed=1):
    rng = np.random.default_rng(seed)
    B, ns, nx = pb["B"], pb["ns"], pb["nx"]
    X = np.zeros((B, ns + 1, nx))
    X[:, :, 0] = rng.uniform(0.02, 1.0, (B, ns + 1))
    X[:, :, 1] = rng.uniform(1.0, 250.0, (B, ns + 1))
    if pb["fat"]:
        pp = np.broadcast_to(pb["params"], (B, fo.NP_FULL))
        X[:, :, 2] = pp[:, None, fo.P_A_REST] * rng.uniform(0.6, 1.0, (B, ns + 1))
        X[:, :, 3] = pp[:, None, fo.P_TAU1_REST] * rng.uniform(1.0, 1.5, (B, ns + 1))
        X[:, :, 4] = pp[:, None, fo.P_KM_REST] * rng.uniform(1.0, 1.5, (B, ns + 1))
    return X


def _bl(t):
    """library [ns, a, b, B] -> oracle [B, ns, a, b]"""
    return np.transpose(t.cpu().numpy(), (3, 0, 1, 2))


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", list(MODELS))
@pytest.mark.parametrize("scheme", [1, 2, 4])
def test_shoot_all_blocks_match_oracle(backend, name, scheme):
    # N = 7 > FES_TILE so that the control / pulse-time tiles need two passes (one partial)
    pb = random_problem(name, B=9, N=7, ns=8, seed=21, shared_ctrl=False)
    S = 4
    X = _X(pb)
    params, _, win, ctrl = to_lib(pb, backend)
    Xl = torch.tensor(np.ascontiguousarray(np.transpose(X, (1, 2, 0))), device=backend)
    want = ["x", "u", "p", "w"] + (["pf"] if pb["fat"] else [])
    res = fes.shoot(pb["kind"], pb["fat"], params, Xl, win, pb["dt"], n_steps=S, ctrl=ctrl, scheme=scheme, want=want,
                    endpoints=True)
    pdir = P_COLS[pb["kind"]] + (PF_COLS if pb["fat"] else [])
    ref = foc.shoot(pb["kind"], pb["fat"], scheme, pb["N"], pb["ns"], S, pb["dt"], 0.0, pb["tauc"], fo.R0_KM_RELATIONSHIP,
                    pb["params"], X, oracle_win(pb), pb["u"], pdir=pdir, want_x=True, want_u=pb["nu"] > 0, want_w=True)
    G = np.transpose(res.G.cpu().numpy(), (2, 0, 1))
    Phi = np.transpose(res.Phi.cpu().numpy(), (2, 0, 1))
    assert rel_err(Phi, X[:, 1:] - ref["G"]) < RTOL64
    assert np.max(np.abs(G - ref["G"])) < RTOL64 * np.max(np.abs(Phi))
    assert entry_err(_bl(res.Jx), ref["Jx"], JKEEP) < RTOL64
    if pb["nu"]:
        assert res.Ju.shape[2] == pb["nu"]
        assert entry_err(_bl(res.Ju), ref["Ju"], JKEEP) < RTOL64
    else:
        assert res.Ju is None
    Jp = _bl(res.Jp)
    assert Jp.shape == ref["Jp"].shape
    assert entry_err(Jp, ref["Jp"], JKEEP) < RTOL64, (name, res.p_columns)
    assert res.Jw.shape[2] == pb["N"]
    assert entry_err(_bl(res.Jw), ref["Jw"], JKEEP) < RTOL64   # pulse-time block: same bar as the others


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", list(MODELS))
def test_shoot_fp32_mode(backend, name):
    """fes_shoot_f32 against the fp64 oracle at north_star's 1e-4: end points per state, every Jacobian entry on its own scale."""
    pb = random_problem(name, B=9, N=7, ns=8, seed=21, shared_ctrl=False)
    S = 10
    X = _X(pb)
    params, _, win, ctrl = to_lib(pb, backend, dtype=torch.float32)
    Xl = torch.tensor(np.ascontiguousarray(np.transpose(X, (1, 2, 0))), dtype=torch.float32, device=backend)
    X32 = np.transpose(Xl.cpu().numpy().astype(np.float64), (2, 0, 1))   # the oracle starts from the same rounded nodes
    want = ["x", "u", "p", "w"] + (["pf"] if pb["fat"] else [])
    res = fes.shoot(pb["kind"], pb["fat"], params, Xl, win, pb["dt"], n_steps=S, ctrl=ctrl, want=want, endpoints=True)
    assert res.Phi.dtype == torch.float32 and res.Jx.dtype == torch.float32
    pdir = P_COLS[pb["kind"]] + (PF_COLS if pb["fat"] else [])
    ref = foc.shoot(pb["kind"], pb["fat"], 4, pb["N"], pb["ns"], S, pb["dt"], 0.0, pb["tauc"], fo.R0_KM_RELATIONSHIP,
                    pb["params"], X32, oracle_win(pb), pb["u"], pdir=pdir, want_x=True, want_u=pb["nu"] > 0, want_w=True)
    Phi = np.transpose(res.Phi.cpu().numpy(), (2, 0, 1))
    assert entry_err(Phi, X32[:, 1:] - ref["G"], keep=(2,)) < RTOL32
    G = np.transpose(res.G.cpu().numpy(), (2, 0, 1))
    assert np.max(np.abs(G - ref["G"]) / np.max(np.abs(Phi), axis=(0, 1))) < RTOL32   # a defect is a difference of states
    for key, J in (("Jx", res.Jx), ("Ju", res.Ju), ("Jp", res.Jp), ("Jw", res.Jw)):
        if J is not None:
            assert entry_err(_bl(J), ref[key], JKEEP) < RTOL32, key


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", ["ding2003", "ding2007_with_fatigue", "hmed2018"])
def test_shoot_defects_only_and_shared_inputs(backend, name):
    pb = random_problem(name, B=33, N=10, ns=5, seed=4, shared_params=True, shared_ctrl=True)
    X = _X(pb)
    params, _, win, ctrl = to_lib(pb, backend)
    Xl = torch.tensor(np.ascontiguousarray(np.transpose(X, (1, 2, 0))), device=backend)
    res = fes.shoot(pb["kind"], pb["fat"], params, Xl, win, pb["dt"], n_steps=10, ctrl=ctrl, want=())
    ref = foc.shoot(pb["kind"], pb["fat"], 4, pb["N"], pb["ns"], 10, pb["dt"], 0.0, pb["tauc"], fo.R0_KM_RELATIONSHIP,
                    pb["params"], X, oracle_win(pb), pb["u"], want_x=False)
    assert res.Jx is None and res.Ju is None and res.Jp is None and res.Jw is None
    scale = np.max(np.abs(X[:, 1:]))
    assert np.max(np.abs(np.transpose(res.G.cpu().numpy(), (2, 0, 1)) - ref["G"])) < RTOL64 * scale
    # u-only request (no x block) goes through the FES_DIR_U kernels
    if pb["nu"]:
        res_u = fes.shoot(pb["kind"], pb["fat"], params, Xl, win, pb["dt"], n_steps=10, ctrl=ctrl, want=("u",))
        ref_u = foc.shoot(pb["kind"], pb["fat"], 4, pb["N"], pb["ns"], 10, pb["dt"], 0.0, pb["tauc"], fo.R0_KM_RELATIONSHIP,
                          pb["params"], X, oracle_win(pb), pb["u"], want_x=False, want_u=True)
        assert entry_err(_bl(res_u.Ju), ref_u["Ju"], JKEEP) < RTOL64


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_shoot_per_trajectory_windows(backend):
    pb = random_problem("hmed2018_with_fatigue", B=6, N=4, ns=5, seed=8, per_traj_win=True, shared_ctrl=False)
    X = _X(pb)
    params, _, win, ctrl = to_lib(pb, backend)
    Xl = torch.tensor(np.ascontiguousarray(np.transpose(X, (1, 2, 0))), device=backend)
    res = fes.shoot(pb["kind"], pb["fat"], params, Xl, win, pb["dt"], n_steps=3, ctrl=ctrl, want=("x", "u", "w"))
    ref = foc.shoot(pb["kind"], pb["fat"], 4, pb["N"], pb["ns"], 3, pb["dt"], 0.0, pb["tauc"], fo.R0_KM_RELATIONSHIP,
                    pb["params"], X, oracle_win(pb), pb["u"], want_x=True, want_u=True, want_w=True)
    assert entry_err(_bl(res.Jx), ref["Jx"], JKEEP) < RTOL64
    assert entry_err(_bl(res.Ju), ref["Ju"], JKEEP) < RTOL64
    assert entry_err(_bl(res.Jw), ref["Jw"], JKEEP) < RTOL64


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_chained_shoot_jacobians_equal_ivp_sensitivities(backend):
    """Size-independent property: the product of interval transition matrices along the single-shooting
    trajectory equals the IVP's chained d x(T)/d x0."""
    pb = random_problem("ding2007_with_fatigue", B=12, N=5, ns=6, seed=13)
    params, x0, win, ctrl = to_lib(pb, backend)
    iv = fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], n_steps=10, ctrl=ctrl, n_shooting=pb["ns"],
                 dirs=fes.FES_DIR_X0, trace="nodes")
    sh = fes.shoot(pb["kind"], pb["fat"], params, iv.trace, win, pb["dt"], n_steps=10, ctrl=ctrl, want=("x",))
    assert torch.max(torch.abs(sh.G)) < 1e-9 * torch.max(torch.abs(iv.trace))   # the IVP nodes are defect-free
    M = torch.eye(pb["nx"], dtype=torch.float64, device=backend)[:, :, None].repeat(1, 1, pb["B"])
    for k in range(pb["ns"]):
        M = torch.einsum("ijb,jkb->ikb", sh.Jx[k], M)
    scale = iv.sens_final.abs().amax(dim=2, keepdim=True)   # every (row, column) entry on its own scale
    assert bool((M[scale.expand_as(M) == 0] == 0).all())     # structural zeros (d Cn / d anything but Cn0) stay zero
    assert float(((M - iv.sens_final).abs() / scale.clamp_min(1e-300)).max()) < 1e-9


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", ["ding2003_with_fatigue", "ding2007_with_fatigue", "hmed2018"])
def test_shoot_plan_equals_one_shot_calls_and_reuses_its_tables(backend, name):
    """ShootPlan = shoot() with the set-up done once: same bits on every call, for new node states / controls / parameters."""
    from cocofest_b200 import _lib

    pb = random_problem(name, B=9, N=7, ns=8, seed=21, shared_ctrl=False)
    params, _, win, ctrl = to_lib(pb, backend)
    want = ("x", "u", "p", "w")
    plan = fes.ShootPlan(pb["kind"], pb["fat"], params, win, pb["dt"], pb["ns"], pb["B"], n_steps=4, ctrl=ctrl, want=want,
                         endpoints=True)
    n0 = _lib.lib().fes_launch_count()
    for it in range(3):
        X = torch.tensor(np.ascontiguousarray(np.transpose(_X(pb, seed=it), (1, 2, 0))), device=backend)
        u_it = None if ctrl is None else (ctrl * (1.0 + 0.01 * it)).contiguous()
        p_it = (params * (1.0 + 0.001 * it)).contiguous()
        got = plan(X, ctrl=u_it, params=p_it)
        ref = fes.shoot(pb["kind"], pb["fat"], p_it, X, win, pb["dt"], n_steps=4, ctrl=u_it, want=want, endpoints=True)
        for a, b in ((got.G, ref.G), (got.Phi, ref.Phi), (got.Jx, ref.Jx), (got.Ju, ref.Ju), (got.Jp, ref.Jp), (got.Jw, ref.Jw)):
            assert (a is None and b is None) or torch.equal(a, b)
        assert got.p_columns == ref.p_columns
    if backend.type == "cuda":   # the plan builds each table layout once; the one-shot call rebuilds it per layout change
        per_plan_call = len(plan._passes)
        assert _lib.lib().fes_launch_count() - n0 >= 3 * per_plan_call
    with pytest.raises(ValueError, match="X: expected"):
        plan(X[:-1].contiguous())
