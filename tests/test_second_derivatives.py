"""Second derivatives of the interval map for an exact-Hessian solver (FesShooting.second_derivatives: central differences of
the exact Jacobian blocks), against an independent route through the pinned numpy oracle: the first derivative by the complex
step (exact to rounding), differenced centrally in the second variable."""
import numpy as np
import pytest
import torch

import cocofest_b200 as cf
from cocofest_b200.ocp_fes import FesShooting
from oracle import fes_oracle as fo
from tests.conftest import BACKENDS


def _setup(name, backend, B=3):
    stim = [k / 20 for k in range(8)]
    model = cf.ModelMaker.create_model(name, stim_time=stim, sum_stim_truncation=5)
    sh = FesShooting(model, 0.4, ode_solver=cf.OdeSolver.RK4(n_integration_steps=5), device=backend)
    rng = np.random.default_rng(4)
    ns, nx = sh.n_shooting, sh.nx
    X = np.zeros((ns + 1, nx, B))
    X[:, 0] = rng.uniform(0.05, 0.8, (ns + 1, B))
    X[:, 1] = rng.uniform(5.0, 150.0, (ns + 1, B))
    if nx == 5:
        rest = model.standard_rest_values()[:, 0]
        X[:, 2] = rest[2] * rng.uniform(0.7, 1.0, (ns + 1, B))
        X[:, 3] = rest[3] * rng.uniform(1.0, 1.3, (ns + 1, B))
        X[:, 4] = rest[4] * rng.uniform(1.0, 1.3, (ns + 1, B))
    U = rng.uniform(3e-4, 6e-4, (ns, 1, B)) if sh.nu == 1 else None
    return model, sh, X, U


def _oracle_second(model, sh, X, U, k, b):
    """[nx, nz, nz] at interval k of scenario b: complex step in z_e, central difference in z_d."""
    nx, nu = sh.nx, sh.nu
    nz = nx + nu
    p = model.param_vector()
    win = sh.window.cpu().numpy()[k]
    z0 = np.concatenate([X[k, :, b], U[k, :, b] if nu else []])
    scale = np.array(list(FesShooting._STATE_SCALE[:nx]) + [1e-30] * nu)

    def first(z):   # d xf / d z_e for all e, by complex step
        J = np.zeros((nx, nz))
        for e in range(nz):
            zc = z.astype(complex)
            zc[e] += 1e-30j
            xall = fo.integrate_interval(model.kind, model.with_fatigue, 4, sh.n_steps, k * sh.dt, sh.dt, zc[:nx],
                                         zc[nx:] if nu else None, p, win, model.tauc)
            J[:, e] = xall[:, -1].imag / 1e-30
        return J

    H = np.zeros((nx, nz, nz))
    for d in range(nz):
        h = 3e-6 * max(abs(z0[d]), scale[d])   # NOT the library's step: truncation errors must not cancel in the comparison
        zp, zm = z0.copy(), z0.copy()
        zp[d] += h
        zm[d] -= h
        H[:, :, d] = (first(zp) - first(zm)) / (zp[d] - zm[d])
    return 0.5 * (H + H.transpose(0, 2, 1))


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", ["ding2003", "ding2007_with_fatigue"])
def test_second_derivatives_match_the_oracle_route(backend, name):
    model, sh, X, U = _setup(name, backend)
    D2 = sh.second_derivatives(torch.tensor(X, device=backend), None if U is None else torch.tensor(U, device=backend))
    ns, nx = sh.n_shooting, sh.nx
    nz = nx + sh.nu
    assert tuple(D2.shape) == (ns, nx, nz, nz, X.shape[-1])
    assert torch.equal(D2, D2.transpose(2, 3))                      # symmetric by construction ...
    got = D2.cpu().numpy()
    for k, b in ((0, 0), (3, 1), (ns - 1, 2)):
        ref = _oracle_second(model, sh, X, U, k, b)
        scale = np.max(np.abs(ref), axis=(1, 2), keepdims=True)     # per output row
        assert np.max(np.abs(got[k, :, :, :, b] - ref) / np.maximum(scale, 1e-300)) < 1e-8, (k, b)   # measured 1e-10
    # ... and the raw difference quotients were symmetric to begin with (mixed partials commute): a size-independent check
    # that the u-perturbed x-block and the x-perturbed u-block agree
    if sh.nu:
        lam = torch.ones((ns, nx, X.shape[-1]), dtype=torch.float64, device=backend)
        H = FesShooting.lagrangian_hessian(lam, D2)
        assert tuple(H.shape) == (ns, nz, nz, X.shape[-1])
        assert float((H - H.transpose(1, 2)).abs().max()) == 0.0
        # Cn decouples from every other state: no second derivative of the Cn row, none with respect to Cn0 except through F
        assert float(D2[:, 0].abs().max()) < 1e-6 * float(D2.abs().max())
