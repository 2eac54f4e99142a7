"""Pins the CPU oracle (oracle/fes_oracle.py and the C restatement) against everything the reference
holds for the path: the golden force trajectories of tests/shard1/test_ivp.py and the RHS / window /
Jacobian fixtures produced by executing the reference's own model classes
(tests/golden/make_reference_fixtures.py)."""
import json
from pathlib import Path

import numpy as np
import pytest

from oracle import fes_oracle as fo
from oracle import fes_oracle_c as foc

GOLD = json.loads((Path(__file__).parent / "golden" / "ivp_golden.json").read_text())
RHS = json.loads((Path(__file__).parent / "golden" / "reference_rhs.json").read_text())

MODELS = {
    "ding2003": (fo.DING2003, False),
    "ding2003_with_fatigue": (fo.DING2003, True),
    "ding2007": (fo.DING2007, False),
    "ding2007_with_fatigue": (fo.DING2007, True),
    "hmed2018": (fo.HMED2018, False),
    "hmed2018_with_fatigue": (fo.HMED2018, True),
}


def golden_problem(name, stim=(0, 0.1, 0.2), final_time=0.3, trunc=3, pulse_mode="single"):
    """The protocol of reference tests/shard1/test_ivp.py:17-124."""
    kind, fat = MODELS[name]
    st = fo.pulse_mode_times(list(stim), pulse_mode)
    ns = fo.prepare_n_shooting(st, final_time)
    win, idx = fo.stim_windows(st, ns, final_time, trunc)
    p = fo.default_params(kind)
    u = None
    if kind == fo.DING2007:
        u = fo.ivp_controls_pulse_width(GOLD["pulse_width"], idx, ns)
    if kind == fo.HMED2018:
        u = fo.ivp_controls_pulse_intensity(GOLD["pulse_intensity"], idx, ns, trunc)
    return dict(kind=kind, fat=fat, ns=ns, dt=final_time / ns, win=win, u=u, p=p, tauc=fo.default_tauc(kind),
                x0=fo.rest_state(kind, fat, p))


@pytest.mark.parametrize("name", list(MODELS))
def test_numpy_oracle_matches_reference_golden_force(name):
    pb = golden_problem(name)
    assert pb["ns"] == 3
    xall = fo.ivp(pb["kind"], pb["fat"], fo.RK4, 10, pb["dt"], pb["x0"], pb["u"], pb["p"], pb["win"], pb["tauc"])
    np.testing.assert_almost_equal(xall[1], np.array(GOLD["F"][name]), decimal=GOLD["decimal"])


@pytest.mark.parametrize("name", list(MODELS))
def test_c_oracle_matches_numpy_oracle_and_golden(name):
    pb = golden_problem(name)
    out = foc.ivp(pb["kind"], pb["fat"], 4, 3, pb["ns"], 10, pb["dt"], pb["tauc"], fo.R0_KM_RELATIONSHIP,
                  pb["p"][None], pb["x0"][None], pb["win"][None], None if pb["u"] is None else pb["u"][None],
                  trace_B=1, trace_every=1)
    np.testing.assert_almost_equal(out["trace"][0, :, 1], np.array(GOLD["F"][name]), decimal=GOLD["decimal"])
    xall = fo.ivp(pb["kind"], pb["fat"], fo.RK4, 10, pb["dt"], pb["x0"], pb["u"], pb["p"], pb["win"], pb["tauc"])
    np.testing.assert_allclose(out["trace"][0].T, xall, rtol=1e-13, atol=1e-300)


def test_pulse_modes_doublet_and_leaked_triplet():
    """test_ivp.py:127-169.  The 'triplet' golden is reproduced only with the doublet-expanded list
    expanded again (the module-level model object is mutated by IvpFes, SURVEY.md section 4 trap 2)."""
    name = "ding2003_with_fatigue"
    pb = golden_problem(name, pulse_mode="single")
    xall = fo.ivp(pb["kind"], True, fo.RK4, 10, pb["dt"], pb["x0"], None, pb["p"], pb["win"], pb["tauc"])
    np.testing.assert_almost_equal(xall[1], np.array(GOLD["F"]["pulse_mode_single"]), decimal=7)
    pb = golden_problem(name, pulse_mode="doublet")
    assert pb["ns"] == 60
    xall = fo.ivp(pb["kind"], True, fo.RK4, 10, pb["dt"], pb["x0"], None, pb["p"], pb["win"], pb["tauc"])
    np.testing.assert_almost_equal(xall[1][150:180], np.array(GOLD["F"]["pulse_mode_doublet_150_180"]), decimal=7)
    leaked = fo.pulse_mode_times(fo.pulse_mode_times([0, 0.1, 0.2], "doublet"), "triplet")
    assert len(leaked) == 18
    ns = fo.prepare_n_shooting(leaked, 0.3)
    win, _ = fo.stim_windows(leaked, ns, 0.3, 3)
    xall = fo.ivp(pb["kind"], True, fo.RK4, 10, 0.3 / ns, pb["x0"], None, pb["p"], win, pb["tauc"])
    np.testing.assert_almost_equal(xall[1][350:380], np.array(GOLD["F"]["pulse_mode_triplet_350_380"]), decimal=7)


@pytest.mark.parametrize("name", list(MODELS))
def test_window_builder_matches_reference_classes(name):
    m = RHS["models"][name]
    prev = [t for t in m["previous_stim"]["time"] if t > -1e6]
    win, idx = fo.stim_windows(m["stim_time"], m["n_shooting"], m["final_time"], m["sum_stim_truncation"], prev)
    np.testing.assert_array_equal(win, np.array(m["window"]))
    assert idx == m["stim_idx_at_node_list"]
    kind, fat = MODELS[name]
    np.testing.assert_array_equal(fo.rest_state(kind, fat, fo.default_params(kind)), np.array(m["rest"]))


def test_literal_float_compare_window_documents_the_trap():
    ref = RHS["literal_window_0p3_over_3"]
    win, idx = fo.stim_windows([0, 0.1, 0.2], 3, 0.3, 3, literal_compare=True)
    np.testing.assert_array_equal(win, np.array(ref["window"]))
    win2, _ = fo.stim_windows([0, 0.1, 0.2], 3, 0.3, 3)
    assert win2[1][-1] == 0.1 and win[1][-1] == 0.0


def _params_from_constants(kind, c):
    p = fo.default_params(kind)
    p[fo.P_A_REST] = c["a_scale"] if kind == fo.DING2007 else c["a_rest"]
    for key, i in (("km_rest", fo.P_KM_REST), ("tau1_rest", fo.P_TAU1_REST), ("tau2", fo.P_TAU2),
                   ("alpha_a", fo.P_ALPHA_A), ("alpha_tau1", fo.P_ALPHA_TAU1), ("alpha_km", fo.P_ALPHA_KM),
                   ("tau_fat", fo.P_TAU_FAT), ("pd0", fo.P_PD0), ("pdt", fo.P_PDT), ("ar", fo.P_AR),
                   ("bs", fo.P_BS), ("Is", fo.P_IS), ("cr", fo.P_CR)):
        if key in c:
            p[i] = c[key]
    return p


PKEY = {"a_rest": fo.P_A_REST, "a_scale": fo.P_A_REST, "km_rest": fo.P_KM_REST, "tau1_rest": fo.P_TAU1_REST,
        "tau2": fo.P_TAU2, "alpha_a": fo.P_ALPHA_A, "alpha_tau1": fo.P_ALPHA_TAU1, "alpha_km": fo.P_ALPHA_KM,
        "tau_fat": fo.P_TAU_FAT, "pd0": fo.P_PD0, "pdt": fo.P_PDT, "ar": fo.P_AR, "bs": fo.P_BS, "Is": fo.P_IS,
        "cr": fo.P_CR}


@pytest.mark.parametrize("name", list(MODELS))
def test_rhs_and_its_jacobians_match_reference_classes(name):
    """xdot and the complex-step Jacobians produced by the reference's system_dynamics."""
    kind, fat = MODELS[name]
    m = RHS["models"][name]
    p = _params_from_constants(kind, m["constants"])
    tauc = m["constants"]["tauc"]
    win_all = np.array(m["window"])
    for pt in m["points"]:
        w = win_all[pt["node"]]
        x = np.array(pt["x"])
        u = np.array(pt["u"]) if pt["u"] else None
        xdot = fo.rhs(kind, fat, pt["t"], x, u, p, w, tauc)
        np.testing.assert_allclose(xdot, np.array(pt["xdot"]), rtol=1e-13, atol=1e-13)
        jx = fo.jacobian_complex_step(lambda z: fo.rhs(kind, fat, pt["t"], z, u, p.astype(complex), w, tauc), x)
        np.testing.assert_allclose(jx, np.array(pt["dxdot_dx_cols"]).T, rtol=1e-12, atol=1e-12)
        if u is not None:
            ju = fo.jacobian_complex_step(lambda z: fo.rhs(kind, fat, pt["t"], x.astype(complex), z, p.astype(complex), w, tauc), u)
            np.testing.assert_allclose(ju, np.array(pt["dxdot_du_cols"]).T, rtol=1e-12, atol=1e-12)
        jw = fo.jacobian_complex_step(lambda z: fo.rhs(kind, fat, pt["t"], x.astype(complex), u, p.astype(complex), z, tauc), w)
        np.testing.assert_allclose(jw, np.array(pt["dxdot_dwin_cols"]).T, rtol=1e-12, atol=1e-10)
        for key, col in pt["dxdot_dp"].items():
            def f(z, key=key):
                pp = p.astype(complex)
                pp[PKEY[key]] = z[0]
                return fo.rhs(kind, fat, pt["t"], x.astype(complex), u, pp, w, tauc)
            jp = fo.jacobian_complex_step(f, np.array([p[PKEY[key]]]))[:, 0]
            np.testing.assert_allclose(jp, np.array(col), rtol=1e-12, atol=1e-12, err_msg=f"{name} d/d{key}")


def test_component_pins():
    """Values asserted by reference tests/shard1/test_models_dynamics_without_bioptim.py that survive."""
    pins = RHS["component_pins"]
    assert np.isclose(np.exp(-(0.1 - 0.09) / 0.02), pins["exp_time_fun(0.1,0.09)"], rtol=1e-15)
    assert np.isclose(1 + (1.05 - 1) * np.exp(-0.1 / 0.02), pins["ri_fun(1.05,0.1)"], rtol=1e-15)
    assert np.isclose(fo.f_dot(5, 100, 3009, 0.050957, 0.103, 0.060), pins["f_dot_fun(5,100,3009,0.050957,0.103)"], rtol=1e-15)
    assert np.isclose(pins["exp_time_fun(0.1,0.09)"], 0.6065306597126332)
    assert np.isclose(pins["f_dot_fun(5,100,3009,0.050957,0.103)"], 2037.0703505791284)


@pytest.mark.parametrize("name", ["ding2003_with_fatigue", "ding2007_with_fatigue", "hmed2018"])
def test_c_oracle_dual_jacobians_match_complex_step_of_pinned_forward_map(name):
    """Derivatives are 'parity unpinned' in the reference; they are established by complex-step
    differentiation of the pinned numpy forward map and must agree with the C dual-number path."""
    kind, fat = MODELS[name]
    rng = np.random.default_rng(7)
    N, ns, S = 4, 3, 5
    dt = 0.05
    stim = [0.0, 0.05, 0.1]
    win, _ = fo.stim_windows(stim, ns, ns * dt, N, previous_time=[-0.07, -0.03])
    p = fo.default_params(kind)
    nx = 5 if fat else 2
    X = np.zeros((ns + 1, nx))
    X[:, 0] = rng.uniform(0.05, 1.0, ns + 1)
    X[:, 1] = rng.uniform(5, 200, ns + 1)
    if fat:
        X[:, 2] = p[fo.P_A_REST] * rng.uniform(0.7, 1, ns + 1)
        X[:, 3] = p[fo.P_TAU1_REST] * rng.uniform(1, 1.3, ns + 1)
        X[:, 4] = p[fo.P_KM_REST] * rng.uniform(1, 1.3, ns + 1)
    nu = {fo.DING2003: 0, fo.DING2007: 1, fo.HMED2018: N}[kind]
    U = None
    if nu == 1:
        U = rng.uniform(2e-4, 6e-4, (ns, 1))
    elif nu:
        U = rng.uniform(30, 120, (ns, nu))
    pdir = [fo.P_A_REST, fo.P_KM_REST, fo.P_TAU1_REST, fo.P_TAU2]
    tauc = fo.default_tauc(kind)
    out = foc.shoot(kind, fat, 4, N, ns, S, dt, 0.0, tauc, fo.R0_KM_RELATIONSHIP, p[None], X[None], win[None],
                    None if U is None else U[None], pdir=pdir, want_x=True, want_u=nu > 0, want_w=True)
    phi, g = fo.shoot(kind, fat, fo.RK4, S, dt, X, U, p, win, tauc)
    np.testing.assert_allclose(out["G"][0], g, rtol=1e-12, atol=1e-12)
    for k in range(ns):
        uk = None if U is None else U[k].astype(complex)
        end = lambda x0, u=uk, pp=p.astype(complex), w=win[k]: fo.integrate_interval(
            kind, fat, fo.RK4, S, dt * k, dt, x0, u, pp, w, tauc)[:, -1]
        jx = fo.jacobian_complex_step(lambda z: end(z), X[k])
        np.testing.assert_allclose(out["Jx"][0, k], jx, rtol=1e-10, atol=1e-12)
        if nu:
            ju = fo.jacobian_complex_step(lambda z: end(X[k].astype(complex), u=z), U[k])
            np.testing.assert_allclose(out["Ju"][0, k], ju, rtol=1e-10, atol=1e-12)
        jw = fo.jacobian_complex_step(lambda z: end(X[k].astype(complex), w=z), win[k])
        np.testing.assert_allclose(out["Jw"][0, k], jw, rtol=1e-10, atol=1e-9)
        cols = []
        for j in pdir:
            def f(z, j=j):
                pp = p.astype(complex)
                pp[j] = z[0]
                return end(X[k].astype(complex), pp=pp)
            cols.append(fo.jacobian_complex_step(f, np.array([p[j]]))[:, 0])
        np.testing.assert_allclose(out["Jp"][0, k], np.stack(cols, axis=1), rtol=1e-10, atol=1e-12)
