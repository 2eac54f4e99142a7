"""The force factor of ``f_dot_fun`` (reference cocofest/models/ding2003.py:309-311): F' = (A q - F/(tau1 + tau2 q)) * (fl * fv + fp).

Outside the FES-MSK models the reference integrates the default arguments (1 * 1 + 0); with them it evaluates fl, fv, fp from the
joint states inside the same ODE.  ``fes.shoot(..., force_factor=c)`` takes c = fl * fv + fp per interval as an INPUT and returns
dPhi/dc beside dPhi/dx and dPhi/du.  Checked against the oracle, which multiplies the F row by the factor exactly where the
reference expression does (C port with dual numbers, and the numpy restatement for the forward map), and against the plain
path at c = 1.  Derivatives are unpinned by the reference: the dual-number oracle is cross-checked here by central
differences of the numpy restatement.
"""
import numpy as np
import pytest
import torch

from cocofest_b200 import fes
from oracle import fes_oracle as fo
from oracle import fes_oracle_c as foc
from tests.conftest import BACKENDS
from tests.problems import MODELS, entry_err, oracle_win, random_problem, rel_err, to_lib
from tests.test_parity_shoot import JKEEP, RTOL32, RTOL64, _X, _bl


def _factor(pb, seed=3, shared=False):
    rng = np.random.default_rng(seed)
    return rng.uniform(0.4, 1.3, (1 if shared else pb["B"], pb["ns"]))   # fl fv + fp of a moving joint: around 1, never 0


def _lib_factor(fc, device, dtype=torch.float64):
    return torch.tensor(fc[0] if fc.shape[0] == 1 else fc.T.copy(), dtype=dtype, device=device)


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", list(MODELS))
@pytest.mark.parametrize("scheme", [1, 2, 4])
def test_shoot_with_force_factor_matches_oracle(backend, name, scheme):
    # N = 7 > FES_TILE: Hmed2018 needs a second intensity tile, which must integrate the same factor (set U | C)
    pb = random_problem(name, B=9, N=7, ns=8, seed=31, shared_ctrl=False)
    S = 5
    X = _X(pb)
    fc = _factor(pb)
    params, _, win, ctrl = to_lib(pb, backend)
    Xl = torch.tensor(np.ascontiguousarray(np.transpose(X, (1, 2, 0))), device=backend)
    res = fes.shoot(pb["kind"], pb["fat"], params, Xl, win, pb["dt"], n_steps=S, ctrl=ctrl, scheme=scheme, endpoints=True,
                    force_factor=_lib_factor(fc, backend))
    ref = foc.shoot(pb["kind"], pb["fat"], scheme, pb["N"], pb["ns"], S, pb["dt"], 0.0, pb["tauc"], fo.R0_KM_RELATIONSHIP,
                    pb["params"], X, oracle_win(pb), pb["u"], want_x=True, want_u=pb["nu"] > 0, force_factor=fc)
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
    Jc = np.transpose(res.Jc.cpu().numpy(), (2, 0, 1))     # [B, ns, nx]
    assert Jc.shape == ref["Jc"].shape
    assert entry_err(Jc, ref["Jc"], (2,)) < RTOL64
    # the numpy restatement of the reference expression, forward map (batch last there)
    un = None if pb["u"] is None else np.transpose(pb["u"], (1, 2, 0))
    phi_np, _ = fo.shoot(pb["kind"], pb["fat"], scheme, S, pb["dt"], np.transpose(X, (1, 2, 0)), un, pb["params"].T,
                         np.transpose(oracle_win(pb), (1, 2, 0)), pb["tauc"], force_factor=fc.T)
    assert rel_err(Phi, np.transpose(phi_np, (2, 0, 1))) < RTOL64


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", ["ding2007_with_fatigue", "hmed2018"])
def test_unit_force_factor_is_the_plain_path_and_a_shared_factor_is_accepted(backend, name):
    pb = random_problem(name, B=6, N=5, ns=4, seed=32)
    X = _X(pb)
    params, _, win, ctrl = to_lib(pb, backend)
    Xl = torch.tensor(np.ascontiguousarray(np.transpose(X, (1, 2, 0))), device=backend)
    plain = fes.shoot(pb["kind"], pb["fat"], params, Xl, win, pb["dt"], ctrl=ctrl, want=("x", "u"))
    one = fes.shoot(pb["kind"], pb["fat"], params, Xl, win, pb["dt"], ctrl=ctrl, force_factor=torch.ones(pb["ns"], dtype=torch.float64, device=backend))
    # x * 1.0 is exact: the same numbers, not merely close ones
    assert torch.equal(one.G, plain.G) and torch.equal(one.Jx, plain.Jx) and torch.equal(one.Ju, plain.Ju)
    # d Phi / d c at c = 1 against central differences of the library's own forward map
    h = 1e-6
    up = fes.shoot(pb["kind"], pb["fat"], params, Xl, win, pb["dt"], ctrl=ctrl, endpoints=True,
                   force_factor=torch.full((pb["ns"],), 1 + h, dtype=torch.float64, device=backend)).Phi
    dn = fes.shoot(pb["kind"], pb["fat"], params, Xl, win, pb["dt"], ctrl=ctrl, endpoints=True,
                   force_factor=torch.full((pb["ns"],), 1 - h, dtype=torch.float64, device=backend)).Phi
    fd = ((up - dn) / (2 * h)).cpu().numpy()
    assert entry_err(one.Jc.cpu().numpy(), fd, (1,)) < 1e-4   # difference quotients of O(1e2) values: ~1e-8 absolute noise


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_force_factor_fp32_mode(backend):
    pb = random_problem("ding2007_with_fatigue", B=16, N=6, ns=6, seed=33, shared_ctrl=False)
    X = _X(pb)
    fc = _factor(pb)
    params, _, win, ctrl = to_lib(pb, backend, torch.float32)
    Xl = torch.tensor(np.ascontiguousarray(np.transpose(X, (1, 2, 0))), dtype=torch.float32, device=backend)
    res = fes.shoot(pb["kind"], pb["fat"], params, Xl, win, pb["dt"], ctrl=ctrl, endpoints=True, force_factor=_lib_factor(fc, backend, torch.float32))
    ref = foc.shoot(pb["kind"], pb["fat"], 4, pb["N"], pb["ns"], 10, pb["dt"], 0.0, pb["tauc"], fo.R0_KM_RELATIONSHIP,
                    pb["params"], X, oracle_win(pb), pb["u"], want_x=True, want_u=True, force_factor=fc)
    Phi = np.transpose(res.Phi.double().cpu().numpy(), (2, 0, 1))
    assert rel_err(Phi, X[:, 1:] - ref["G"]) < RTOL32
    assert entry_err(np.transpose(res.Jc.double().cpu().numpy(), (2, 0, 1)), ref["Jc"], (2,)) < RTOL32
    assert entry_err(_bl(res.Jx.double()), ref["Jx"], JKEEP) < RTOL32


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", ["ding2003_with_fatigue", "ding2007_with_fatigue", "hmed2018_with_fatigue"])
def test_shoot_plan_with_force_factor_is_the_one_shot_call(backend, name):
    pb = random_problem(name, B=12, N=7, ns=5, seed=36, shared_ctrl=False)
    X = _X(pb)
    params, _, win, ctrl = to_lib(pb, backend)
    Xl = torch.tensor(np.ascontiguousarray(np.transpose(X, (1, 2, 0))), device=backend)
    plan = None
    for rep in range(2):   # second call: new factors and states through the same plan (tables reused)
        fc = _lib_factor(_factor(pb, seed=40 + rep), backend)
        Xr = (Xl * (1.0 - 0.05 * rep)).contiguous()
        one = fes.shoot(pb["kind"], pb["fat"], params, Xr, win, pb["dt"], ctrl=ctrl, endpoints=True, force_factor=fc)
        rows = ([ctrl[:pb["ns"]]] if ctrl is not None else []) + [fc[:, None, :]]
        ext = torch.cat(rows, dim=1).contiguous()
        if plan is None:
            plan = fes.ShootPlan(pb["kind"], pb["fat"], params, win, pb["dt"], pb["ns"], pb["B"], ctrl=ext, endpoints=True, force_factor=True)
        got = plan(Xr, ctrl=ext)
        assert torch.equal(got.G, one.G) and torch.equal(got.Phi, one.Phi) and torch.equal(got.Jx, one.Jx) and torch.equal(got.Jc, one.Jc)
        assert (got.Ju is None and one.Ju is None) or torch.equal(got.Ju, one.Ju)


def test_oracle_force_factor_column_against_differences_of_the_numpy_restatement():
    """Pins the oracle's own d/dc (dual numbers in C) to the forward map written after the reference expression."""
    pb = random_problem("hmed2018_with_fatigue", B=3, N=4, ns=3, seed=34, shared_ctrl=False)
    X = _X(pb)
    fc = _factor(pb)
    ref = foc.shoot(pb["kind"], pb["fat"], 4, pb["N"], pb["ns"], 6, pb["dt"], 0.0, pb["tauc"], fo.R0_KM_RELATIONSHIP,
                    pb["params"], X, oracle_win(pb), pb["u"], want_x=False, force_factor=fc)
    args = (pb["kind"], pb["fat"], 4, 6, pb["dt"], np.transpose(X, (1, 2, 0)), np.transpose(pb["u"], (1, 2, 0)), pb["params"].T,
            np.transpose(oracle_win(pb), (1, 2, 0)), pb["tauc"])
    h = 1e-6
    up, _ = fo.shoot(*args, force_factor=fc.T * (1 + h))
    dn, _ = fo.shoot(*args, force_factor=fc.T * (1 - h))
    fd = np.transpose((up - dn) / (2 * h * fc.T[:, None, :]), (2, 0, 1))
    assert entry_err(ref["Jc"], fd, (2,)) < 1e-4   # difference quotients: ~1e-8 absolute noise on entries down to 1e-5


def test_force_factor_argument_checks():
    pb = random_problem("ding2007", B=4, N=5, ns=4, seed=35)
    with pytest.raises(ValueError, match="Jx, Ju and Jc"):
        fes.shoot(pb["kind"], pb["fat"], torch.tensor(pb["params"].T.copy()), torch.zeros((5, 2, 4), dtype=torch.float64),
                  torch.tensor(pb["win"]), pb["dt"], ctrl=torch.tensor(pb["u"][0]), want=("x", "p"), force_factor=torch.ones(4, dtype=torch.float64))
    from tests.conftest import _load_emul
    from cocofest_b200 import _lib
    import ctypes

    L = _load_emul()
    desc = fes._desc(fes.FES_DING2007, False, 4, 5, 4, 10, fes.FES_DIR_X0 | fes.FES_DIR_C, 0, 4, 0.05, 0.0, None, fes.R0_KM_RELATIONSHIP)
    buf = _lib.FesShootBuffers(None, None, None, None, None, None, None, 0, 0, None, 0)
    assert L.fes_shoot_f64(ctypes.byref(desc), ctypes.byref(buf), None) == -4      # X0 | C alone exists for Ding2003 only
    assert L.fes_num_dirs(fes.FES_DING2007, 1, fes.FES_DIR_X0 | fes.FES_DIR_U | _lib.FES_DIR_C) == 5 + 1 + 1
    assert L.fes_num_dirs(fes.FES_DING2003, 0, fes.FES_DIR_X0 | fes.FES_DIR_U | _lib.FES_DIR_C) == 2 + 1
    assert L.fes_num_dirs(fes.FES_HMED2018, 1, fes.FES_DIR_U | _lib.FES_DIR_C) == 5 + 1
    assert L.fes_num_dirs(fes.FES_DING2007, 1, fes.FES_DIR_U | _lib.FES_DIR_C) < 0
    desc.dirs = fes.FES_DIR_X0 | fes.FES_DIR_U | _lib.FES_DIR_C
    assert L.fes_ivp_f64(ctypes.byref(desc), ctypes.byref(_lib.FesIvpBuffers()), None) == -4   # no force factor in the ivp entry points
