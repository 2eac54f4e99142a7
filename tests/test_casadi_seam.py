"""`as_casadi_function` and `GpuRK4` (cocofest_b200/bioptim_seam.py) EXECUTED against stand-ins of the casadi Callback
contract and of bioptim's `OdeSolver.RK4` (tests/standins: casadi and bioptim are not installable here).  Checked: the
declarations casadi reads at construct time (n_in / n_out, names, sparsities), evaluation through the Function call,
the Jacobian Function's signature (`jac_<out>_<in>`, outputs outermost, 8 inputs / 12 outputs), which blocks are dense
and which structurally zero, the block values against `FesIntegrator.jacobian`, `.map`, and the object bioptim builds
with `nlp.ode_solver.integrator(ode, ode_opt)` (ode_solver.py:232-271; callers penalty.py:1092-1099)."""
import sys
from pathlib import Path

import numpy as np
import pytest

import cocofest_b200 as cf
from cocofest_b200.bioptim_seam import FesIntegrator, GpuRK4, as_casadi_function
from tests.conftest import BACKENDS
from tests.test_bioptim_seam import ID_PARAMS, _case

STANDINS = str(Path(__file__).resolve().parent / "standins")


@pytest.fixture
def standins(monkeypatch):
    for mod in ("casadi", "bioptim"):
        if mod in sys.modules and "standin" not in str(getattr(sys.modules[mod], "__file__", "")):
            pytest.skip(f"the real {mod} is installed: the maintainers' own suite exercises the wrappers")
    monkeypatch.syspath_prepend(STANDINS)
    for mod in ("casadi", "bioptim"):
        monkeypatch.delitem(sys.modules, mod, raising=False)
    import casadi

    assert "standin" in casadi.__version__
    yield casadi
    for mod in ("casadi", "bioptim"):
        sys.modules.pop(mod, None)


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", ["ding2003", "ding2007_with_fatigue", "hmed2018", "hmed2018_with_fatigue"])
def test_callback_pair_declarations_values_and_jacobian_blocks(backend, standins, name):
    casadi = standins
    S = 5
    model, t_span, x0, u, p, d, names = _case(name, 1, seed=7)
    it = FesIntegrator(model, cf.OdeSolver.RK4(n_integration_steps=S), parameter_names=names, device=backend)
    nx, nu, npar, N = it.nx, it.nu, len(names), it.N
    f = as_casadi_function(it, name="integrator")
    # what casadi read at construct time
    assert (f.n_in(), f.n_out()) == (6, 2)
    assert f.name_in() == ["t_span", "x0", "u", "p", "a", "d"] and f.name_out() == ["xf", "xall"]     # integrator.py:134
    assert [f.size_in(i) for i in range(6)] == [(2, 1), (nx, 1), (nu, 1), (npar, 1), (0, 1), (N, 1)]
    assert [f.size_out(i) for i in range(2)] == [(nx, 1), (nx, S + 1)]
    args = [t_span, x0[:, 0], None if u is None else u[:, 0], p[:, 0], [], d]
    xf, xall = f(*args)
    ref = it(t_span, x0[:, 0], None if u is None else u[:, 0], p[:, 0], None, d)
    np.testing.assert_array_equal(np.array(xf)[:, 0], ref["xf"][:, 0].cpu().numpy())
    np.testing.assert_array_equal(np.array(xall), ref["xall"][..., 0].cpu().numpy())
    # the Jacobian Function casadi asks for: inputs = 6 inputs + the 2 nominal outputs, outputs = jac_<out>_<in>
    J = f.jacobian()
    assert (J.n_in(), J.n_out()) == (8, 12)
    assert J._names_out_expected[:6] == ["jac_xf_t_span", "jac_xf_x0", "jac_xf_u", "jac_xf_p", "jac_xf_a", "jac_xf_d"]
    assert J._names_out_expected[6] == "jac_xall_t_span"
    sizes_in = [2, nx, nu, npar, 0, N]
    for o in range(2):
        for k in range(6):
            sp = J.sparsity_out(o * 6 + k)
            assert sp.shape() == ((nx if o == 0 else nx * (S + 1)), sizes_in[k])
            dense = o == 0 and k in (1, 2, 3)
            assert sp.nnz() == (sp.numel() if dense else 0), (o, k)      # t_span, a, d and every xall block: structurally zero
    blocks = J(*args, xf, xall)
    want = it.jacobian(t_span, x0[:, 0], None if u is None else u[:, 0], p[:, 0], None, d)
    np.testing.assert_array_equal(np.array(blocks[1]), want["x0"][..., 0].cpu().numpy())
    if nu:
        np.testing.assert_array_equal(np.array(blocks[2]), want["u"][..., 0].cpu().numpy())
    else:
        assert blocks[2].shape == (nx, 0)
    np.testing.assert_array_equal(np.array(blocks[3]), want["p"][..., 0].cpu().numpy())
    for i in (0, 4, 5, 6, 7, 8, 9, 10, 11):
        assert not np.array(blocks[i]).any()
    # the dense Jacobian in casadi's input order is the same set of blocks
    D = it.dense_jacobian(t_span, x0[:, 0], None if u is None else u[:, 0], p[:, 0], None, d)
    np.testing.assert_array_equal(D[:, 2:2 + nx], np.array(blocks[1]))
    np.testing.assert_array_equal(D[:, 2 + nx + nu:2 + nx + nu + npar], np.array(blocks[3]))
    # casadi refuses arguments of the wrong shape before eval is reached
    with pytest.raises(RuntimeError, match="dimension mismatch"):
        f(t_span, x0[:-1, 0], args[2], args[3], [], d)


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_gpu_rk4_is_what_bioptim_instantiates_per_phase(backend, standins):
    import bioptim

    name, S, B = "ding2007_with_fatigue", 4, 3
    model, t_span, x0, u, p, d, names = _case(name, B, seed=9)
    solver = GpuRK4(n_integration_steps=S, parameter_names=names, device=backend)
    # cocofest's own check on the solver type (fes_ocp.py:334-338) and bioptim's read of the step count
    assert isinstance(solver, bioptim.OdeSolver.RK4) and not isinstance(solver, bioptim.OdeSolver.RK2)
    assert solver.n_integration_steps == S
    integ = solver.initialize_integrator({}, {"model": model})      # nlp.ode_solver.integrator(ode, ode_opt)
    assert integ.shape_xf == (5, 1) and integ.shape_xall == (5, S + 1)
    np.testing.assert_allclose(integ.step_times_from_dt(t_span), t_span[0] + t_span[1] * np.arange(S + 1) / S)
    assert integ.function.name_in() == ["t_span", "x0", "u", "p", "a", "d"]
    xf, xall = integ(t_span, x0[:, 0], u[:, 0], p[:, 0], [], d)      # state_continuity: nlp.dynamics[k](...)["xf"]
    core = integ.core(t_span, x0, u, p, None, d)
    np.testing.assert_array_equal(np.array(xf)[:, 0], core["xf"][:, 0].cpu().numpy())
    # penalty_option.py:596-600 maps the function over the nodes: columns side by side, t_span / d shared here
    mapped = integ.map(B, "thread", 2)
    xf_m, xall_m = mapped(t_span, x0, u, p, np.zeros((0, B)), d)
    np.testing.assert_array_equal(np.array(xf_m), core["xf"].cpu().numpy())
    assert xall_m.shape == (5, B * (S + 1))
    np.testing.assert_array_equal(np.array(xall_m)[:, (S + 1):2 * (S + 1)], core["xall"][..., 1].cpu().numpy())


def test_wrappers_still_refuse_without_the_modules():
    """Without casadi / bioptim (this image, no stand-in on the path) nothing is evaluated in their place."""
    for mod in ("casadi", "bioptim"):
        if mod in sys.modules:
            pytest.skip(f"{mod} importable")
    model = cf.DingModelFrequency(stim_time=[0, 0.1], sum_stim_truncation=2)
    it = FesIntegrator(model, cf.OdeSolver.RK4(5), device="cpu")
    with pytest.raises(ImportError, match="casadi"):
        as_casadi_function(it)
    with pytest.raises(ImportError, match="bioptim"):
        GpuRK4(10)
