"""Force-tracking target and objective of OcpFes (fourier_approx.py:7-38, fes_ocp.py:530-571): the Fourier fit
against values produced by the reference's own class (tests/golden/make_fourier_fixture.py), the batched cost and
its gradient against a direct restatement and central differences."""
import json
from pathlib import Path

import numpy as np
import pytest
import torch

from cocofest_b200.fourier_approx import FourierSeries, force_to_track, tracking_objective

GOLD = json.loads((Path(__file__).parent / "golden" / "fourier_golden.json").read_text())


@pytest.mark.parametrize("case", GOLD["cases"], ids=lambda c: f"n{c['n']}_ns{c['n_shooting']}")
def test_fourier_fit_matches_reference_class(case):
    x, y = np.array(GOLD["x"]), np.array(GOLD["y"])
    fs = FourierSeries()
    ab = fs.compute_real_fourier_coeffs(x, y, case["n"])
    np.testing.assert_allclose(ab, np.array(case["ab"]), rtol=1e-12, atol=1e-12)
    nodes = np.linspace(0, 1, case["n_shooting"] + 1)
    val = fs.fit_func_by_fourier_series_with_real_coeffs(nodes, ab)
    np.testing.assert_allclose(val, np.array(case["force_to_track"]), rtol=1e-12, atol=1e-10)
    if case["n"] == 50:
        np.testing.assert_allclose(force_to_track([x, y], case["n_shooting"]), np.array(case["force_to_track"]),
                                   rtol=1e-12, atol=1e-10)
    # fourier_approx (fourier_approx.py:40-49, without its plot) = the two calls above at the sample points
    np.testing.assert_array_equal(fs.fourier_approx(x, y, case["n"]), fs.fit_func_by_fourier_series_with_real_coeffs(x, ab))
    assert float(fs.fit_func_by_fourier_series_with_real_coeffs(0.25, ab)) == pytest.approx(
        fs.fit_func_by_fourier_series_with_real_coeffs(np.array([0.25]), ab)[0])


def test_tracking_objective_cost_and_gradient():
    rng = np.random.default_rng(0)
    ns, nx, B, dt = 12, 5, 7, 1.0 / 33
    X = torch.tensor(rng.uniform(0, 200, (ns + 1, nx, B)))
    target = rng.uniform(0, 200, ns + 1)
    end = rng.uniform(50, 150, B)
    cost, grad = tracking_objective(X, dt, target=target, end_node_target=end)
    F = X[:, 1].numpy()
    ref = 100 * dt * ((F - target[:, None]) ** 2).sum(axis=0) + (F[-1] - end) ** 2
    np.testing.assert_allclose(cost.numpy(), ref, rtol=1e-13)
    assert not grad[:, [0, 2, 3, 4]].any()
    for k in (0, 5, ns):
        Xp, Xm = X.clone(), X.clone()
        Xp[k, 1] += 1e-4
        Xm[k, 1] -= 1e-4
        fd = (tracking_objective(Xp, dt, target, end)[0] - tracking_objective(Xm, dt, target, end)[0]) / 2e-4
        np.testing.assert_allclose(grad[k, 1].numpy(), fd.numpy(), rtol=1e-7)
    c2, g2 = tracking_objective(X, dt, target=np.tile(target[:, None], (1, B)))
    c3, _ = tracking_objective(X, dt, end_node_target=end)
    np.testing.assert_allclose((c2 + c3).numpy(), ref, rtol=1e-13)
    with pytest.raises(ValueError, match="rows"):
        tracking_objective(X, dt, target=target[:-1])


def test_muscle_level_objectives_cost_and_gradient():
    """best fit / overall fatigue / overall force production (custom_objectives.py:43-117 as fes_ocp_dynamics.py:675-693 weights
    them): closed forms and gradients against central differences."""
    from cocofest_b200.fourier_approx import best_fit_objective, muscle_fatigue_objective, muscle_force_objective

    rng = np.random.default_rng(1)
    M, ns, nx, B, dt = 3, 6, 5, 4, 0.05
    X = torch.tensor(rng.uniform(20, 200, (M, ns + 1, nx, B)))
    X[:, :, 2] = torch.tensor(rng.uniform(2000, 3000, (M, ns + 1, B)))
    a_rest = np.array([3009.0, 2800.0, 3100.0])
    target = rng.uniform(20, 200, ns + 1)

    def fd(fun, T, eps=1e-6):
        g = torch.zeros_like(T)
        flat, gf = T.reshape(-1), g.reshape(-1)
        for i in range(flat.numel()):
            old = float(flat[i])
            h = eps * max(abs(old), 1.0)
            flat[i] = old + h
            cp = fun(T)
            flat[i] = old - h
            cm = fun(T)
            flat[i] = old
            gf[i] = float((cp - cm).sum()) / (2 * h)      # each entry moves one scenario's cost only
        return g

    c, g = muscle_force_objective(X, dt)
    np.testing.assert_allclose(c.numpy(), dt * (X[:, :, 1].numpy() ** 2).sum(axis=(0, 1)), rtol=1e-14)
    np.testing.assert_allclose(g.numpy(), fd(lambda T: muscle_force_objective(T, dt)[0], X.clone()).numpy(), rtol=1e-6, atol=1e-6)
    Xe = X[:, -1].clone()
    c, g = muscle_fatigue_objective(Xe, a_rest)
    np.testing.assert_allclose(c.numpy(), (a_rest.sum() / Xe[:, 2].numpy().sum(axis=0)) ** 2, rtol=1e-14)
    np.testing.assert_allclose(g.numpy(), fd(lambda T: muscle_fatigue_objective(T, a_rest)[0], Xe.clone()).numpy(), rtol=1e-6, atol=1e-12)
    c1, g1 = muscle_fatigue_objective(Xe[0], a_rest[:1])          # one muscle: [nx, B]
    np.testing.assert_allclose(c1.numpy(), (a_rest[0] / Xe[0, 2].numpy()) ** 2, rtol=1e-14)
    assert g1.shape == Xe[0].shape
    X1 = X[0].clone()
    c, g = best_fit_objective(X1, dt, target)
    np.testing.assert_allclose(c.numpy(), dt * ((1 - target[:, None] / X1[:, 1].numpy()) ** 2).sum(axis=0), rtol=1e-14)
    np.testing.assert_allclose(g.numpy(), fd(lambda T: best_fit_objective(T, dt, target)[0], X1.clone()).numpy(), rtol=1e-6, atol=1e-9)


def test_custom_objective_namespace_matches_the_functions():
    """cocofest_b200.CustomObjective: the reference's method names on batched node tensors."""
    import cocofest_b200 as cf

    x, y = np.array(GOLD["x"]), np.array(GOLD["y"])
    ab = FourierSeries().compute_real_fourier_coeffs(x, y, 50)
    rng = np.random.default_rng(2)
    ns, B, dt = 20, 3, 0.05
    X = torch.tensor(rng.uniform(20, 200, (ns + 1, 5, B)))
    c, g = cf.CustomObjective.track_state_from_time(X, dt, ab, key="F")
    c_ref, g_ref = tracking_objective(X, dt, target=force_to_track([x, y], ns))
    assert torch.equal(c, c_ref) and torch.equal(g, g_ref)
    force = rng.uniform(20, 200, ns + 1)
    c, _ = cf.CustomObjective.track_state_from_time_interpolate(X, dt, force, "F", "best fit")
    np.testing.assert_allclose(c.numpy(), dt * ((1 - force[:, None] / X[:, 1].numpy()) ** 2).sum(axis=0), rtol=1e-14)
    with pytest.raises(RuntimeError, match="not implemented"):
        cf.CustomObjective.track_state_from_time_interpolate(X, dt, force, "F", "other")
    c, g = cf.CustomObjective.minimize_overall_muscle_force_production(X, dt)
    assert g.shape == X.shape and c.shape == (B,)
    c, g = cf.CustomObjective.minimize_overall_muscle_fatigue(X[-1], [3009.0])
    assert g.shape == X[-1].shape and c.shape == (B,)
    assert callable(cf.GpuRK4) and cf.SolutionToPickle is not None
