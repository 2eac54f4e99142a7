"""``FesIntegrator``: the call contract of bioptim's ``nlp.dynamics[k]`` (dynamics/integrator.py:82-134;
callers penalty.py:1092-1099, solve_ivp_interface.py:67-69) evaluated by the kernels, against the oracle's
restatement of the same interval map (numpy for xall, C dual numbers for the Jacobian blocks)."""
import numpy as np
import pytest
import torch

import cocofest_b200 as cf
from cocofest_b200.bioptim_seam import FesIntegrator
from oracle import fes_oracle as fo
from oracle import fes_oracle_c as foc
from tests.conftest import BACKENDS
from tests.problems import MODELS, rel_err

RTOL64 = 1e-9
NAMES = list(MODELS)
ID_PARAMS = {
    "ding2003": ["a_rest", "km_rest", "tau1_rest", "tau2"],                       # ding2003_force_parameter_identification.py:121-150
    "ding2003_with_fatigue": ["tau2", "a_rest", "alpha_a", "tau_fat"],            # declared order is the caller's
    "ding2007": ["tau1_rest", "tau2", "km_rest", "a_scale", "pd0", "pdt"],        # ding2007_...:105-148
    "ding2007_with_fatigue": ["pd0", "a_scale", "alpha_km"],
    "hmed2018": ["a_rest", "km_rest", "tau1_rest", "tau2", "ar", "bs", "Is", "cr"],  # hmed2018_...:103-160
    "hmed2018_with_fatigue": ["Is", "cr", "alpha_tau1"],
}
CANON = {n: i for i, n in enumerate(["a_rest", "km_rest", "tau1_rest", "tau2", "alpha_a", "alpha_tau1", "alpha_km",
                                      "tau_fat", "pd0", "pdt", "ar", "bs", "Is", "cr"])}
CANON["a_scale"] = 0


def _case(name, B, seed):
    rng = np.random.default_rng(seed)
    N = 6
    stim = [0.0, 0.05, 0.1, 0.15, 0.2, 0.25, 0.3]
    prev = {"time": [-0.12, -0.05]}
    if name.startswith("hmed"):
        prev["pulse_intensity"] = [40, 45]  # hmed2018.py:43-47 requires one intensity per previous pulse
    model = cf.ModelMaker.create_model(name, stim_time=stim, previous_stim=prev, sum_stim_truncation=N)
    ns, T = 7, 0.35
    data, _ = model.get_numerical_data_time_series(ns, T)
    k = 5                                  # node 5: the window holds 6 real pulses
    d = np.asarray(data["stim_time"][:, 0, k], dtype=float)
    t_span = np.array([k * T / ns, T / ns])
    nx = model.nb_state
    x0 = np.zeros((nx, B))
    x0[0] = rng.uniform(0.05, 0.8, B)
    x0[1] = rng.uniform(1.0, 150.0, B)
    if nx == 5:
        rest = model.standard_rest_values()[:, 0]
        x0[2] = rest[2] * rng.uniform(0.7, 1.0, B)
        x0[3] = rest[3] * rng.uniform(1.0, 1.3, B)
        x0[4] = rest[4] * rng.uniform(1.0, 1.3, B)
    nu = {fo.DING2007: 1, fo.HMED2018: N}.get(model.kind, 0)
    u = None
    if nu:
        u = rng.uniform(2e-4, 6e-4, (nu, B)) if model.kind == fo.DING2007 else rng.uniform(30.0, 120.0, (nu, B))
    names = ID_PARAMS[name]
    base = model.param_vector()
    p = np.stack([base[CANON[n]] * rng.uniform(0.8, 1.2, B) for n in names])
    return model, t_span, x0, u, p, d, names


def _oracle_params(model, names, p, B):
    full = np.tile(model.param_vector(), (B, 1))
    for j, n in enumerate(names):
        full[:, CANON[n]] = p[j]
    return full


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("scheme", [1, 2, 4])
def test_integrator_call_and_jacobian_match_oracle(backend, name, scheme):
    B, S = 4, 5
    model, t_span, x0, u, p, d, names = _case(name, B, seed=11)
    solver = {1: cf.OdeSolver.RK1, 2: cf.OdeSolver.RK2, 4: cf.OdeSolver.RK4}[scheme](n_integration_steps=S)
    it = FesIntegrator(model, solver, parameter_names=names, device=backend)
    assert it.shape_xf == (model.nb_state, 1) and it.shape_xall == (model.nb_state, S + 1)
    np.testing.assert_allclose(it.step_times_from_dt(t_span), t_span[0] + t_span[1] * np.arange(S + 1) / S)
    out = it(t_span, x0, u, p, None, d)
    full = _oracle_params(model, names, p, B)
    for b in range(B):
        ref = fo.integrate_interval(model.kind, model.with_fatigue, scheme, S, t_span[0], t_span[1], x0[:, b],
                                    None if u is None else u[:, b], full[b], d, model.tauc)
        assert rel_err(out["xall"][:, :, b].cpu().numpy(), ref) < RTOL64
        assert rel_err(out["xf"][:, b].cpu().numpy(), ref[:, -1]) < RTOL64
    # Jacobian blocks: C oracle (dual numbers through the reference formulation), one interval
    J = it.jacobian(t_span, x0, u, p, None, d, stim_time=True)
    X = np.stack([x0.T, x0.T], axis=1)
    win = np.stack([d, d])[None]
    ref = foc.shoot(model.kind, model.with_fatigue, scheme, it.N, 1, S, t_span[1], t_span[0], model.tauc,
                    fo.R0_KM_RELATIONSHIP, full, X, win, None if u is None else u.T[:, None, :],
                    pdir=[CANON[n] for n in names], want_x=True, want_u=u is not None, want_w=True)
    assert rel_err(J["xf"].cpu().numpy().T, x0.T - ref["G"][:, 0]) < RTOL64
    assert rel_err(np.transpose(J["x0"].cpu().numpy(), (2, 0, 1)), ref["Jx"][:, 0]) < RTOL64
    if u is not None:
        assert J["u"].shape[1] == it.nu
        assert rel_err(np.transpose(J["u"].cpu().numpy(), (2, 0, 1)), ref["Ju"][:, 0]) < RTOL64
    else:
        assert J["u"] is None
    Jp = np.transpose(J["p"].cpu().numpy(), (2, 0, 1))
    assert Jp.shape == ref["Jp"][:, 0].shape
    for c, n in enumerate(names):
        assert rel_err(Jp[..., c], ref["Jp"][:, 0, :, c]) < RTOL64, n
    assert rel_err(np.transpose(J["d"].cpu().numpy(), (2, 0, 1)), ref["Jw"][:, 0], floor=1e-9) < 10 * RTOL64


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_single_column_like_the_reference_call_and_dense_layout(backend):
    """One scenario, vectors in, as bioptim calls it; the dense Jacobian has casadi's input ordering."""
    model, t_span, x0, u, p, d, names = _case("ding2007_with_fatigue", 1, seed=3)
    it = FesIntegrator(model, cf.OdeSolver.RK4(n_integration_steps=10), parameter_names=names, device=backend)
    out = it(t_span, x0[:, 0], u[:, 0], p[:, 0], [], d)
    assert tuple(out["xf"].shape) == (5, 1) and tuple(out["xall"].shape) == (5, 11, 1)
    assert torch.equal(out["xall"][:, -1], out["xf"]) and np.allclose(out["xall"][:, 0, 0].cpu().numpy(), x0[:, 0])
    D = it.dense_jacobian(t_span, x0[:, 0], u[:, 0], p[:, 0], [], d)
    assert D.shape == (5, 2 + 5 + 1 + len(names) + it.N)
    assert not D[:, :2].any() and not D[:, 2 + 5 + 1 + len(names):].any()
    # central differences of the integrator's own xf (layout check, not the precision test)
    z = np.concatenate([x0[:, 0], u[:, 0], p[:, 0]])
    for c in range(z.size):
        hstep = 1e-4 * abs(z[c])
        zp, zm = z.copy(), z.copy()
        zp[c] += hstep
        zm[c] -= hstep
        f = lambda v: it(t_span, v[:5], v[5:6], v[6:], [], d)["xf"][:, 0].cpu().numpy()
        fd = (f(zp) - f(zm)) / (2 * hstep)
        noise = 1e-14 * np.abs(f(z)) / hstep  # rounding of xf itself, amplified by the difference quotient
        assert np.all(np.abs(fd - D[:, 2 + c]) <= 1e-5 * np.abs(D[:, 2 + c]) + noise), c


@pytest.mark.parametrize("backend", [BACKENDS[0]], indirect=True)
def test_argument_checks(backend):
    model, t_span, x0, u, p, d, names = _case("hmed2018", 2, seed=5)
    with pytest.raises(ValueError, match="not identifiable"):
        FesIntegrator(model, cf.OdeSolver.RK4(5), parameter_names=["alpha_a"], device=backend)
    with pytest.raises(ValueError, match="ode_solver"):
        FesIntegrator(model, object(), device=backend)
    it = FesIntegrator(model, cf.OdeSolver.RK4(5), parameter_names=names, device=backend)
    with pytest.raises(ValueError, match="x0 must be"):
        it(t_span, x0[:1], u, p, None, d)
    with pytest.raises(ValueError, match="u must have"):
        it(t_span, x0, None, p, None, d)
    with pytest.raises(ValueError, match="columns"):
        it(t_span, x0, u[:, :1].repeat(3, axis=1), p, None, d)
