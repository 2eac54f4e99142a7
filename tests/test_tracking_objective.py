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
