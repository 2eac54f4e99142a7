"""cocofest_b200.hill_coefficients against values produced by the reference's own functions
(tests/golden/make_hill_fixture.py -> hill_golden.json) and its derivatives against central differences."""
import json
from pathlib import Path

import numpy as np
import torch

from cocofest_b200 import hill_coefficients as hc

GOLD = json.loads((Path(__file__).parent / "golden" / "hill_golden.json").read_text())


def test_relationships_match_the_reference_functions():
    rows = GOLD["rows"]
    nl = np.array([r["norm_length"] for r in rows])
    v = np.array([r["velocity"] for r in rows])
    assert np.allclose(hc.muscle_force_length_coefficient(nl), [r["fl"] for r in rows], rtol=1e-14, atol=0)
    assert np.allclose(hc.muscle_force_velocity_coefficient(v), [r["fv"] for r in rows], rtol=1e-13, atol=1e-15)
    assert np.allclose(hc.muscle_passive_force_coefficient(nl), [r["fp"] for r in rows], rtol=1e-14, atol=0)
    ref = np.array([r["fl"] * r["fv"] + r["fp"] for r in rows])
    assert np.allclose(hc.force_factor(nl, v), ref, rtol=1e-13)
    # torch tensors go through the same code
    got = hc.force_factor(torch.tensor(nl), torch.tensor(v))
    assert isinstance(got, torch.Tensor) and np.allclose(got.numpy(), ref, rtol=1e-13)
    # switched-off relationships contribute 1, 1, 0 (the default arguments of f_dot_fun, ding2003.py:281-283)
    assert hc.force_factor(force_length=False, force_velocity=False, passive_force=False) == 1.0
    assert np.allclose(hc.force_factor(nl, None, force_velocity=False), np.array([r["fl"] + r["fp"] for r in rows]), rtol=1e-13)


def test_known_points_of_the_curves():
    assert abs(hc.muscle_force_length_coefficient(np.array(1.0)) - 1.0) < 5e-3      # normalised: ~1 at the optimal length
    assert abs(hc.muscle_force_velocity_coefficient(np.array(0.0)) - 1.0) < 5e-3    # isometric
    assert hc.muscle_passive_force_coefficient(np.array(0.9)) == 0.0                # clamped below the optimal length
    assert abs(hc.muscle_passive_force_coefficient(np.array(1.6)) - 1.0) < 1e-12    # (exp(kpe) - 1)/(exp(kpe) - 1) at 1 + e0


def test_derivatives_against_central_differences():
    nl = np.linspace(0.5, 1.6, 47)   # a grid that does not hit the kink of the passive clamp at the optimal length
    v = np.linspace(-5, 5, 47)
    h = 1e-6
    c, dl, dv = hc.force_factor(nl, v, derivative=True)
    fd_l = (hc.force_factor(nl + h, v) - hc.force_factor(nl - h, v)) / (2 * h)
    fd_v = (hc.force_factor(nl, v + h) - hc.force_factor(nl, v - h)) / (2 * h)
    assert np.allclose(dl, fd_l, rtol=1e-6, atol=1e-8) and np.allclose(dv, fd_v, rtol=1e-6, atol=1e-8)
    assert np.allclose(c, hc.force_factor(nl, v))
