"""Generates golden fixtures by executing the REFERENCE'S OWN model classes in the build container.

    python tests/golden/make_reference_fixtures.py      (needs /root/reference; writes reference_rhs.json)

casadi / bioptim are not installable here, and the reference's model files import them at module top.
The model arithmetic itself only needs ``exp``, ``tanh`` and ``vertcat``; this script therefore installs
two stand-in modules in ``sys.modules`` -- ``casadi`` (exp/tanh -> numpy, vertcat -> np.array, MX/SX as
empty type names) and ``bioptim`` (every attribute an inert class, never called by the code we run) --
and imports ``cocofest/models/*.py`` straight from /root/reference without executing
``cocofest/__init__.py``.  Every number written below is produced by the reference's
``system_dynamics`` / ``get_numerical_data_time_series`` / ``min_pulse_intensity`` code, evaluated
with IEEE doubles (which is also what casadi's SX virtual machine does).  Because numpy's exp/tanh accept
complex input, the same reference code also yields complex-step Jacobians of its right-hand side, which
is the only derivative information obtainable from the reference in this container.
"""
import importlib
import json
import sys
import types
from pathlib import Path

import numpy as np

REF = "/root/reference"
OUT = Path(__file__).with_name("reference_rhs.json")


def _install_stubs():
    class _Stub(types.ModuleType):
        def __getattr__(self, name):
            if name.startswith("__"):
                raise AttributeError(name)
            return type(name, (object,), {"__init__": lambda self, *a, **k: None})

    cas = _Stub("casadi")
    cas.exp, cas.tanh = np.exp, np.tanh
    cas.vertcat = lambda *a: np.array(a)
    cas.MX, cas.SX = type("MX", (), {}), type("SX", (), {})
    sys.modules["casadi"] = cas
    sys.modules["bioptim"] = _Stub("bioptim")
    pkg = types.ModuleType("cocofest")
    pkg.__path__ = [REF + "/cocofest"]
    sys.modules["cocofest"] = pkg
    sub = types.ModuleType("cocofest.models")
    sub.__path__ = [REF + "/cocofest/models"]
    sys.modules["cocofest.models"] = sub


def main():
    _install_stubs()
    mods = {
        "ding2003": ("ding2003", "DingModelFrequency"),
        "ding2003_with_fatigue": ("ding2003_with_fatigue", "DingModelFrequencyWithFatigue"),
        "ding2007": ("ding2007", "DingModelPulseWidthFrequency"),
        "ding2007_with_fatigue": ("ding2007_with_fatigue", "DingModelPulseWidthFrequencyWithFatigue"),
        "hmed2018": ("hmed2018", "DingModelPulseIntensityFrequency"),
        "hmed2018_with_fatigue": ("hmed2018_with_fatigue", "DingModelPulseIntensityFrequencyWithFatigue"),
    }
    rng = np.random.default_rng(20251104)
    fixture = {"how": "reference model classes executed with numpy stand-ins for casadi.exp/tanh/vertcat", "models": {}}

    for name, (mod, cls) in mods.items():
        klass = getattr(importlib.import_module(f"cocofest.models.{mod}"), cls)
        n_trunc = 5
        stim = [0.0, 0.05, 0.1, 0.15, 0.2, 0.25, 0.3, 0.35]
        prev = {"time": [-0.1, -0.05]}
        if "hmed" in name:
            prev["pulse_intensity"] = [45, 55]
        model = klass(stim_time=list(stim), previous_stim=prev, sum_stim_truncation=n_trunc)
        fatigue = bool(model._with_fatigue)
        ns, final_time = 8, 0.4
        d, stim_idx = model.get_numerical_data_time_series(ns, final_time)
        win = d["stim_time"][:, 0, :].T  # [ns+1, N]
        entry = {
            "sum_stim_truncation": n_trunc,
            "stim_time": stim,
            "previous_stim": {k: [float(v) for v in vals] for k, vals in model.previous_stim.items()},
            "n_shooting": ns,
            "final_time": final_time,
            "window": win.tolist(),
            "stim_idx_at_node_list": [[int(i) for i in row] for row in stim_idx],
            "rest": np.asarray(model.standard_rest_values(), dtype=float).ravel().tolist(),
            "constants": {
                k: float(getattr(model, k))
                for k in ("tauc", "r0_km_relationship", "a_rest", "tau1_rest", "tau2", "km_rest", "alpha_a",
                          "alpha_tau1", "alpha_km", "tau_fat", "a_scale", "pd0", "pdt", "ar", "bs", "Is", "cr")
                if hasattr(model, k)
            },
            "points": [],
        }
        if "hmed" in name:
            entry["min_pulse_intensity"] = float(model.min_pulse_intensity())

        def call(x, t, w, u):
            kw = dict(cn=x[0], f=x[1], t=t, t_stim_prev=w)
            if fatigue:
                kw.update(a=x[2], tau1=x[3], km=x[4])
            if "ding2007" in name:
                kw["pulse_width"] = u[0]
            if "hmed" in name:
                kw["pulse_intensity"] = u
            return np.asarray(model.system_dynamics(**kw))

        for node in (1, 3, 4, 7):
            w = win[node]
            for _ in range(4):
                t = node * final_time / ns + rng.uniform(0, final_time / ns)
                x = [rng.uniform(0.01, 1.5), rng.uniform(0.0, 300.0)]
                if fatigue:
                    a0 = model.a_scale if "ding2007" in name else model.a_rest
                    x += [a0 * rng.uniform(0.6, 1.0), model.tau1_rest * rng.uniform(1, 1.6),
                          model.km_rest * rng.uniform(1, 1.6)]
                if "ding2007" in name:
                    u = [rng.uniform(2e-4, 6e-4)]
                elif "hmed" in name:
                    u = list(rng.uniform(30, 120, size=n_trunc))
                else:
                    u = []
                xdot = call(np.array(x), t, w, np.array(u))
                # complex-step Jacobians of the reference RHS w.r.t. states and controls
                eps = 1e-30
                jx = []
                for j in range(len(x)):
                    xc = np.array(x, dtype=complex)
                    xc[j] += 1j * eps
                    jx.append((np.imag(call(xc, t, w, np.array(u))) / eps).tolist())
                ju = []
                for j in range(len(u)):
                    uc = np.array(u, dtype=complex)
                    uc[j] += 1j * eps
                    ju.append((np.imag(call(np.array(x), t, w, uc)) / eps).tolist())
                # ... and w.r.t. the window times (pulse-time sensitivities)
                jw = []
                for j in range(n_trunc):
                    wc = np.array(w, dtype=complex)
                    wc[j] += 1j * eps
                    jw.append((np.imag(call(np.array(x), t, wc, np.array(u))) / eps).tolist())
                # ... and w.r.t. the identifiable parameters (set through the reference's own attributes)
                jp = {}
                for key in model.identifiable_parameters:
                    old = getattr(model, key)
                    setattr(model, key, old + 1j * eps)
                    jp[key] = (np.imag(call(np.array(x), t, w, np.array(u))) / eps).tolist()
                    setattr(model, key, old)
                entry["points"].append(
                    {"node": node, "t": float(t), "x": [float(v) for v in x], "u": [float(v) for v in u],
                     "xdot": np.real(xdot).tolist(), "dxdot_dx_cols": jx, "dxdot_du_cols": ju,
                     "dxdot_dwin_cols": jw, "dxdot_dp": jp}
                )
        fixture["models"][name] = entry

    # the literal float-compare window of the golden-vector protocol (documents SURVEY §4 trap 1)
    klass = getattr(importlib.import_module("cocofest.models.ding2003"), "DingModelFrequency")
    m = klass(stim_time=[0, 0.1, 0.2], sum_stim_truncation=3)
    d, idx = m.get_numerical_data_time_series(3, 0.3)
    fixture["literal_window_0p3_over_3"] = {"window": d["stim_time"][:, 0, :].T.tolist(), "stim_idx_at_node_list": idx}
    # component pins quoted by tests/shard1/test_models_dynamics_without_bioptim.py
    m = klass()
    fixture["component_pins"] = {
        "exp_time_fun(0.1,0.09)": float(m.exp_time_fun(0.1, 0.09)),
        "ri_fun(1.05,0.1)": float(m.ri_fun(1.05, 0.1)),
        "f_dot_fun(5,100,3009,0.050957,0.103)": float(m.f_dot_fun(5, 100, 3009, 0.050957, 0.103)),
        "cn_dot_fun(0,5)": float(m.cn_dot_fun(0, 5)),
    }
    OUT.write_text(json.dumps(fixture, indent=1))
    print("wrote", OUT, OUT.stat().st_size, "bytes")


if __name__ == "__main__":
    main()
