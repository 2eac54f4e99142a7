"""The reference-facing Python surface: model classes, ModelMaker, IvpFes (same dictionaries, defaults and
error messages as cocofest/integration/ivp_fes.py; cases follow reference tests/shard1/test_ivp.py), OcpFes
grid helper, FesShooting Jacobian layout; plus: the C-ABI library exports every symbol of include/fes_b200.h."""
import ctypes
import json
import re
from pathlib import Path

import numpy as np
import pytest
import torch

import cocofest_b200 as cf
from cocofest_b200 import _lib, fes
from tests.conftest import BACKENDS, ROOT

GOLD = json.loads((Path(__file__).parent / "golden" / "ivp_golden.json").read_text())


def test_library_exports_every_symbol_of_the_header():
    header = (ROOT / "include" / "fes_b200.h").read_text()
    declared = set(re.findall(r"\b(fes_[a-z0-9_]+)\s*\(", header))
    assert declared == set(_lib.PROTOTYPES), declared ^ set(_lib.PROTOTYPES)
    so = _lib.LIB_PATH
    assert so.exists(), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    import subprocess
    syms = subprocess.run(["nm", "-D", "--defined-only", str(so)], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T (fes_[a-z0-9_]+)", syms))
    assert declared <= exported, declared - exported
    # struct layouts match the header (sizes the C compiler gives)
    assert ctypes.sizeof(_lib.FesDesc) == 12 * 4 + 2 * 8 + 4 * 8
    assert ctypes.sizeof(_lib.FesIvpBuffers) == 10 * 8
    assert ctypes.sizeof(_lib.FesShootBuffers) == 7 * 8 + 2 * 4 + 2 * 8


def test_product_fails_loudly_without_the_cuda_library(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, "_LIB", None)
    monkeypatch.setattr(_lib, "LIB_PATH", tmp_path / "libfes_b200.so")
    with pytest.raises(_lib.FesLibraryError, match="no CPU fallback"):
        _lib.lib()


def test_model_maker_and_model_attributes():
    names = ["ding2003", "ding2003_with_fatigue", "ding2007", "ding2007_with_fatigue", "hmed2018", "hmed2018_with_fatigue"]
    for n in names:
        m = cf.ModelMaker.create_model(n, stim_time=[0, 0.1], sum_stim_truncation=3)
        assert isinstance(m, cf.FesModel)
        assert m.nb_state == (5 if "fatigue" in n else 2)
        assert m.name_dof == ["Cn", "F", "A", "Tau1", "Km"][: m.nb_state]
        assert m.standard_rest_values().shape == (m.nb_state, 1)
        cls, kw = m.serialize()
        assert cls is type(m) and "tauc" in kw
        assert m.param_vector().shape == (fes.FES_NP,)
    with pytest.raises(ValueError, match="Unknown model type"):
        cf.ModelMaker.create_model("ding1999")
    m = cf.DingModelFrequencyWithFatigue()
    assert (m.alpha_a, m.alpha_tau1, m.alpha_km, m.tau_fat) == (-4.0 * 10e-2, 2.1 * 10e-6, 1.9 * 10e-6, 127)
    m.set_a_rest(None, 3500.0)
    m.set_tau_fat(None, 100.0)
    p = m.param_vector()
    assert p[0] == 3500.0 and p[7] == 100.0
    m7 = cf.DingModelPulseWidthFrequencyWithFatigue(sum_stim_truncation=5)
    assert m7.tauc == 0.011 and m7.standard_rest_values()[2, 0] == 4920 and m7.param_vector()[0] == 4920
    assert set(m7.identifiable_parameters) == {"a_scale", "tau1_rest", "km_rest", "tau2", "pd0", "pdt", "alpha_a",
                                               "alpha_tau1", "alpha_km", "tau_fat"}
    mh = cf.DingModelPulseIntensityFrequency()
    assert np.isclose(mh.min_pulse_intensity(), 17.02854931878943)
    assert np.isclose(mh.lambda_i_calculation(30), 0.0799499, atol=1e-6)
    with pytest.raises(ValueError, match="same length"):
        cf.DingModelPulseIntensityFrequency(previous_stim={"time": [-0.1], "pulse_intensity": []})


def test_scalar_helpers_match_reference_component_pins():
    """Values from reference tests/shard1/test_models_dynamics_without_bioptim.py that are still valid."""
    m = cf.DingModelFrequency()
    assert np.isclose(m.exp_time_fun(0.1, 0.09), 0.6065306597126332)
    assert np.isclose(m.ri_fun(1.05, 0.1), 1.0003368973499542)
    assert np.isclose(m.f_dot_fun(5, 100, 3009, 0.050957, 0.103), 2037.0703505791284)
    assert np.isclose(m.cn_dot_fun(0, 5), 250.0)
    fixtures = json.loads((Path(__file__).parent / "golden" / "reference_rhs.json").read_text())["models"]
    for name, fx in fixtures.items():
        prev = {"time": [t for t in fx["previous_stim"]["time"]]}
        if "pulse_intensity" in fx["previous_stim"]:
            prev["pulse_intensity"] = list(fx["previous_stim"]["pulse_intensity"])
        m = cf.ModelMaker.create_model(name, stim_time=fx["stim_time"], previous_stim=prev, sum_stim_truncation=fx["sum_stim_truncation"])
        data, idx = m.get_numerical_data_time_series(fx["n_shooting"], fx["final_time"])
        np.testing.assert_array_equal(data["stim_time"][:, 0, :].T, np.array(fx["window"]))
        assert idx == fx["stim_idx_at_node_list"]
        for pt in fx["points"]:
            x = pt["x"]
            kw = dict(cn=x[0], f=x[1], t=pt["t"], t_stim_prev=fx["window"][pt["node"]])
            if m.with_fatigue:
                kw.update(a=x[2], tau1=x[3], km=x[4])
            if "ding2007" in name:
                kw["pulse_width"] = pt["u"][0]
            if "hmed" in name:
                kw["pulse_intensity"] = pt["u"]
            np.testing.assert_allclose(m.system_dynamics(**kw), pt["xdot"], rtol=1e-13, atol=1e-13)


# every method name the reference's model classes define (cocofest/models/{fes_model,ding2003*,ding2007*,hmed2018*}.py)
REFERENCE_MODEL_METHODS = {
    "ding2003": ["set_a_rest", "set_km_rest", "set_tau1_rest", "set_tau2", "standard_rest_values", "serialize", "get_r0", "get_lambda_i",
                 "system_dynamics", "exp_time_fun", "ri_fun", "cn_sum_fun", "cn_dot_fun", "f_dot_fun", "calculate_cn_dot", "dynamics",
                 "declare_ding_variables", "get_numerical_data_time_series"],
    "ding2003_with_fatigue": ["set_alpha_a", "set_alpha_tau1", "set_alpha_km", "set_tau_fat", "a_dot_fun", "tau1_dot_fun", "km_dot_fun"],
    "ding2007": ["set_a_scale", "set_pd0", "set_pdt", "a_calculation", "a_calculation_identification", "set_impulse_width"],
    "hmed2018": ["set_ar", "set_bs", "set_Is", "set_cr", "lambda_i_calculation", "lambda_i_calculation_identification",
                 "set_impulse_intensity", "min_pulse_intensity"],
}


def test_every_method_of_the_reference_models_exists():
    for name, methods in REFERENCE_MODEL_METHODS.items():
        m = cf.ModelMaker.create_model(name, stim_time=[0.0, 0.1], sum_stim_truncation=3)
        missing = [fn for fn in methods if not callable(getattr(m, fn, None))]
        assert not missing, (name, missing)


def test_identification_helpers_and_the_dynamics_hook():
    """a_calculation_identification / lambda_i_calculation_identification (custom_constraints.py:41, :93) are the model's own
    expressions with the parameters passed in; the pin 0.0799499 is the reference's
    (tests/shard1/test_models_dynamics_without_bioptim.py:249).  ``dynamics`` is the reference's dyn_func on numbers."""
    m7 = cf.ModelMaker.create_model("ding2007_with_fatigue", stim_time=[0.0, 0.1], sum_stim_truncation=3)
    assert m7.a_calculation_identification(4920, 3e-4, m7.pd0, m7.pdt) == m7.a_calculation(4920, 3e-4)
    m7.set_impulse_width([2e-4, 3e-4])
    assert m7.pulse_width == [2e-4, 3e-4]
    mh = cf.ModelMaker.create_model("hmed2018", stim_time=[0.0, 0.1], sum_stim_truncation=3)
    np.testing.assert_almost_equal(mh.lambda_i_calculation(30), 0.0799499)
    assert mh.lambda_i_calculation_identification(30, mh.ar, mh.bs, mh.Is, mh.cr) == mh.lambda_i_calculation(30)
    mh.set_impulse_intensity(np.array([40.0, 55.0]))
    assert mh.impulse_intensity == [40.0, 55.0]
    fixtures = json.loads((Path(__file__).parent / "golden" / "reference_rhs.json").read_text())["models"]
    for name, fx in fixtures.items():
        prev = {"time": list(fx["previous_stim"]["time"])}
        if "pulse_intensity" in fx["previous_stim"]:
            prev["pulse_intensity"] = list(fx["previous_stim"]["pulse_intensity"])
        m = cf.ModelMaker.create_model(name, stim_time=fx["stim_time"], previous_stim=prev, sum_stim_truncation=fx["sum_stim_truncation"])
        for pt in fx["points"]:
            ev = type(m).dynamics(pt["t"], np.array(pt["x"]), np.array(pt["u"]) if pt.get("u") else None, None, None,
                                  np.array(fx["window"][pt["node"]]), nlp=None, fes_model=m)
            np.testing.assert_allclose(ev.dxdt, pt["xdot"], rtol=1e-13, atol=1e-13)
            assert ev.defects is None
        with pytest.raises(NotImplementedError, match="GpuRK4"):
            m.declare_ding_variables(None, None)


def test_prepare_n_shooting():
    assert cf.OcpFes.prepare_n_shooting([0, 0.1, 0.2], 0.3) == 3
    assert cf.OcpFes.prepare_n_shooting([0, 0.005, 0.1, 0.105, 0.2, 0.205], 0.3) == 60
    assert cf.OcpFes.prepare_n_shooting([s + k / 33 for s in range(0, 4, 2) for k in range(33)], 4) == 132
    assert cf.OcpFes.prepare_n_shooting(list(np.round([i / 33 for i in range(33)], 3)), 1) == 1000


def test_ivp_fes_constructor_errors_follow_the_reference():
    model = cf.DingModelFrequencyWithFatigue(stim_time=[0, 0.1, 0.2], sum_stim_truncation=3)
    with pytest.raises(TypeError, match="model must be a FesModel type"):
        cf.IvpFes({"model": None}, {"final_time": 0.3}, device="cpu")
    with pytest.raises(ValueError, match="final_time must be an int or float type"):
        cf.IvpFes({"model": model}, {"final_time": "0.3"}, device="cpu")
    with pytest.raises(ValueError, match="pulse_mode must be a string type"):
        cf.IvpFes({"model": model, "pulse_mode": 1}, {"final_time": 0.3}, device="cpu")
    with pytest.raises(ValueError, match="ode_solver must be a OdeSolver type"):
        cf.IvpFes({"model": model}, {"final_time": 0.3, "ode_solver": None}, device="cpu")
    with pytest.raises(ValueError, match="n_thread must be a int type"):
        cf.IvpFes({"model": model}, {"final_time": 0.3, "n_threads": "1"}, device="cpu")
    with pytest.raises(ValueError, match="Pulse mode not yet implemented"):
        cf.IvpFes({"model": model, "pulse_mode": "quadruplet"}, {"final_time": 0.3}, device="cpu")
    m7 = cf.DingModelPulseWidthFrequency(stim_time=[0, 0.1, 0.2], sum_stim_truncation=3)
    with pytest.raises(ValueError, match="pulse width must be greater than minimum pulse width"):
        cf.IvpFes({"model": m7, "pulse_width": 0.0001}, {"final_time": 0.3}, device="cpu")
    with pytest.raises(TypeError, match="pulse_width must be int, float or list type"):
        cf.IvpFes({"model": m7, "pulse_width": True}, {"final_time": 0.3}, device="cpu")
    mh = cf.DingModelPulseIntensityFrequency(stim_time=[0, 0.1, 0.2], sum_stim_truncation=3)
    with pytest.raises(ValueError, match="Pulse intensity must be greater than minimum pulse intensity"):
        cf.IvpFes({"model": mh, "pulse_intensity": 0.1}, {"final_time": 0.3}, device="cpu")
    with pytest.raises(ValueError, match="Frequency must be an int"):
        cf.IvpFes.from_frequency_and_final_time({"model": model, "frequency": 10.5, "round_down": True}, {"final_time": 1})
    with pytest.raises(ValueError, match="n_stim must be an int"):
        cf.IvpFes.from_frequency_and_n_stim({"model": model, "frequency": 10, "n_stim": 3.5}, {})


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name,extra", [
    ("ding2003", {}), ("ding2003_with_fatigue", {}),
    ("ding2007", {"pulse_width": [0.0003, 0.0004, 0.0005]}), ("ding2007_with_fatigue", {"pulse_width": [0.0003, 0.0004, 0.0005]}),
    ("hmed2018", {"pulse_intensity": [50, 60, 70]}), ("hmed2018_with_fatigue", {"pulse_intensity": [50, 60, 70]}),
])
def test_ivp_fes_integrate_reproduces_reference_goldens(backend, name, extra):
    """The exact calls of reference tests/shard1/test_ivp.py:17-124, through the mirrored IvpFes."""
    model = cf.ModelMaker.create_model(name, stim_time=[0, 0.1, 0.2], sum_stim_truncation=3)
    fes_parameters = {"model": model, **extra}
    ivp_parameters = {"final_time": 0.3, "ode_solver": cf.OdeSolver.RK4(n_integration_steps=10)}
    ivp = cf.IvpFes(fes_parameters, ivp_parameters, device=backend)
    result, time = ivp.integrate()
    assert set(result) == set(model.name_dof)
    assert result["F"].shape == (1, 31) and time.shape == (31,)
    np.testing.assert_almost_equal(result["F"][0], np.array(GOLD["F"][name]), decimal=7)
    np.testing.assert_allclose(time, np.linspace(0, 0.3, 31), atol=1e-15)


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_ivp_fes_pulse_modes_and_frequency_constructors(backend):
    model = cf.DingModelFrequencyWithFatigue(stim_time=[0, 0.1, 0.2], sum_stim_truncation=3)
    ivp = cf.IvpFes({"model": model, "pulse_mode": "doublet"}, {"final_time": 0.3}, device=backend)
    assert ivp.n_shooting == 60 and model.stim_time == [0, 0.005, 0.1, 0.105, 0.2, 0.205]
    result, _ = ivp.integrate()
    np.testing.assert_almost_equal(result["F"][0][150:180], np.array(GOLD["F"]["pulse_mode_doublet_150_180"]), decimal=7)
    # the reference's "triplet" golden comes from re-using the (mutated) model object (SURVEY.md section 4 trap 2)
    ivp = cf.IvpFes({"model": model, "pulse_mode": "triplet"}, {"final_time": 0.3}, device=backend)
    result, _ = ivp.integrate()
    np.testing.assert_almost_equal(result["F"][0][350:380], np.array(GOLD["F"]["pulse_mode_triplet_350_380"]), decimal=7)
    # README example (config 1): 10 Hz for 1 s, truncation 20 (defaults)
    model = cf.DingModelFrequencyWithFatigue(stim_time=[i / 10 for i in range(10)])
    ivp = cf.IvpFes.from_frequency_and_final_time({"model": model, "frequency": 10, "round_down": True},
                                                  {"final_time": 1}, device=backend)
    assert ivp.n_shooting == 10
    result, time = ivp.integrate()
    assert result["F"].shape == (1, 101) and time.shape == (101,)
    from oracle import fes_oracle as fo
    w, _ = fo.stim_windows([i / 10 for i in range(10)], 10, 1, 20)
    p0 = fo.default_params(fo.DING2003)
    ref = fo.ivp(fo.DING2003, True, fo.RK4, 10, 0.1, fo.rest_state(fo.DING2003, True, p0), None, p0, w, 0.020)
    for i, key in enumerate(model.name_dof):
        np.testing.assert_allclose(result[key][0], ref[i], rtol=1e-9, atol=1e-12)
    # batched extension: B parameter sets of the same protocol, sensitivities w.r.t. the identifiable parameters
    B = 6
    p = torch.tensor(model.param_vector())[:, None].repeat(1, B)
    p[0] *= torch.linspace(0.8, 1.2, B, dtype=torch.float64)
    res = ivp.integrate_batch(params=p, dirs=fes.FES_DIR_P, trace="nodes")
    assert res.trace.shape == (11, 5, B) and res.sens_final.shape == (5, 4, B)
    assert np.isclose(res.trace[-1, 1, B // 2 if False else 0].item(), res.x_final[1, 0].item())
    assert torch.all(res.x_final[1, 1:] > res.x_final[1, :-1])  # force grows with a_rest


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_fes_shooting_jacobian_layout_matches_finite_differences(backend):
    """d g / d v in bioptim's variable ordering, checked by central differences of the defects."""
    model = cf.DingModelPulseWidthFrequencyWithFatigue(stim_time=[0, 0.05, 0.1, 0.15], sum_stim_truncation=3)
    sh = cf.FesShooting(model, final_time=0.2, ode_solver=cf.OdeSolver.RK4(n_integration_steps=5), device=backend)
    ns, nx, nu = sh.n_shooting, sh.nx, sh.nu
    assert (ns, nx, nu) == (4, 5, 1)
    rng = np.random.default_rng(3)
    B = 2
    X = np.zeros((ns + 1, nx, B))
    X[:, 0], X[:, 1] = rng.uniform(0.05, 0.8, (ns + 1, B)), rng.uniform(5, 100, (ns + 1, B))
    X[:, 2], X[:, 3], X[:, 4] = 4920 * rng.uniform(0.8, 1, (ns + 1, B)), 0.0606 * rng.uniform(1, 1.2, (ns + 1, B)), 0.137 * rng.uniform(1, 1.2, (ns + 1, B))
    U = rng.uniform(2e-4, 6e-4, (ns, nu, B))
    res = sh.evaluate(X, U, want=("x", "u", "p"))
    lay = sh.variable_layout(n_params=res.Jp.shape[2])
    assert lay == {"dt": 0, "x": 1, "u": 1 + (ns + 1) * nx, "p": 1 + (ns + 1) * nx + ns * nu, "size": 1 + (ns + 1) * nx + ns * nu + 6}
    Jd = sh.jacobian_dense(res, scenario=1).cpu().numpy()
    assert Jd.shape == (ns * nx, lay["size"])

    def g_of(Xv, Uv, pv):
        out = sh.defects(Xv, Uv, params=pv)
        return np.transpose(out.cpu().numpy(), (2, 0, 1)).reshape(B, -1)[1]

    p0 = model.param_vector()
    pmat = np.repeat(p0[:, None], B, axis=1)
    # state columns
    for k, j in [(0, 0), (1, 1), (2, 2), (3, 3), (4, 4), (2, 0)]:
        h = 1e-6 * max(abs(X[k, j, 1]), 1e-3)
        Xp, Xm = X.copy(), X.copy()
        Xp[k, j, 1] += h
        Xm[k, j, 1] -= h
        fd = (g_of(Xp, U, pmat) - g_of(Xm, U, pmat)) / (2 * h)
        np.testing.assert_allclose(Jd[:, lay["x"] + k * nx + j], fd, rtol=2e-6, atol=1e-6 * np.max(np.abs(fd)))
    for k in range(ns):
        h = 1e-9
        Up, Um = U.copy(), U.copy()
        Up[k, 0, 1] += h
        Um[k, 0, 1] -= h
        fd = (g_of(X, Up, pmat) - g_of(X, Um, pmat)) / (2 * h)
        np.testing.assert_allclose(Jd[:, lay["u"] + k], fd, rtol=1e-5, atol=1e-6 * np.max(np.abs(fd)))
    for c, key in enumerate(res.p_columns):
        i = _lib.PARAM_INDEX[key]
        h = 1e-6 * abs(p0[i])
        pp, pm = pmat.copy(), pmat.copy()
        pp[i, 1] += h
        pm[i, 1] -= h
        fd = (g_of(X, U, pp) - g_of(X, U, pm)) / (2 * h)
        # rows of the A state are O(1e3): central differences carry ~1e-5 absolute noise there
        np.testing.assert_allclose(Jd[:, lay["p"] + c], fd, rtol=1e-4, atol=1e-4 * np.max(np.abs(fd)) + 2e-5, err_msg=key)


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_hmed_sliding_window_controls(backend):
    """Controls built from per-pulse intensities follow custom_constraints.py:101-119 (right-aligned window,
    missing slots = minimum intensity) and reproduce IvpFes' own controls where those are defined."""
    model = cf.DingModelPulseIntensityFrequency(stim_time=[0, 0.1, 0.2], sum_stim_truncation=3)
    sh = cf.FesShooting(model, final_time=0.3, ode_solver=cf.OdeSolver.RK4(10), device=backend)
    U = sh.pulse_intensity_controls([50.0, 60.0, 70.0]).cpu().numpy()
    lo = model.min_pulse_intensity()
    np.testing.assert_allclose(U, [[lo, lo, 50], [lo, 50, 60], [50, 60, 70]])
