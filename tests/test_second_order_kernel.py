"""The second-order tangent kernel (fes_shoot_hess_f64: hyper-dual arithmetic through the RK stages, one thread per
(scenario, interval, pair of directions)) against two independent routes:
  * the pinned numpy oracle differentiated by the complex step (first derivative, exact to rounding) and centrally
    differenced in the second variable (tests/test_second_derivatives.py::_oracle_second);
  * central differences of the library's own exact Jacobian blocks (FesShooting.second_derivatives(method="fd")).
The reference pins no derivative at all: "unpinned by the reference", established here."""
import ctypes

import numpy as np
import pytest
import torch

import cocofest_b200 as cf
from cocofest_b200 import _lib, fes
from cocofest_b200.ocp_fes import FesShooting
from tests.conftest import BACKENDS
from oracle import fes_oracle as fo

MODELS = ["ding2003", "ding2003_with_fatigue", "ding2007", "ding2007_with_fatigue", "hmed2018", "hmed2018_with_fatigue"]


def _setup(name, backend, B=3, scheme="RK4", seed=4):
    stim = [k / 20 for k in range(8)]
    model = cf.ModelMaker.create_model(name, stim_time=stim, sum_stim_truncation=4)
    solver = getattr(cf.OdeSolver, scheme)(n_integration_steps=5)
    sh = FesShooting(model, 0.4, ode_solver=solver, device=backend)
    rng = np.random.default_rng(seed)
    ns, nx = sh.n_shooting, sh.nx
    X = np.zeros((ns + 1, nx, B))
    X[:, 0] = rng.uniform(0.05, 0.8, (ns + 1, B))
    X[:, 1] = rng.uniform(5.0, 150.0, (ns + 1, B))
    if nx == 5:
        rest = model.standard_rest_values()[:, 0]
        X[:, 2] = rest[2] * rng.uniform(0.7, 1.0, (ns + 1, B))
        X[:, 3] = rest[3] * rng.uniform(1.0, 1.3, (ns + 1, B))
        X[:, 4] = rest[4] * rng.uniform(1.0, 1.3, (ns + 1, B))
    U = None
    if sh.nu == 1:
        U = rng.uniform(3e-4, 6e-4, (ns, 1, B))
    elif sh.nu > 1:
        U = rng.uniform(30.0, 110.0, (ns, sh.nu, B))
    return model, sh, X, U


def _oracle_second(model, sh, X, U, k, b):
    """[nx, nz, nz] at interval k of scenario b: complex step in z_e (exact to rounding), Richardson-extrapolated central
    difference in z_d (steps 1e-3 and 5e-4 of the magnitude: truncation O(h^4) ~ 1e-12, rounding ~ 1e-13)."""
    nx, nu = sh.nx, sh.nu
    nz = nx + nu
    p = model.param_vector()
    win = sh.window.cpu().numpy()[k]
    z0 = np.concatenate([X[k, :, b], U[k, :, b] if nu else []])
    scale = np.array(list(FesShooting._STATE_SCALE[:nx]) + [1e-30] * nu)

    def first(z):
        J = np.zeros((nx, nz))
        for e in range(nz):
            zc = z.astype(complex)
            zc[e] += 1e-30j
            xall = fo.integrate_interval(model.kind, model.with_fatigue, sh.scheme, sh.n_steps, k * sh.dt, sh.dt, zc[:nx],
                                         zc[nx:] if nu else None, p, win, model.tauc)
            J[:, e] = xall[:, -1].imag / 1e-30
        return J

    def central(d, h):
        zp, zm = z0.copy(), z0.copy()
        zp[d] += h
        zm[d] -= h
        return (first(zp) - first(zm)) / (zp[d] - zm[d])

    H = np.zeros((nx, nz, nz))
    noise = np.zeros((nx, nz, nz))      # rounding of the complex-step Jacobian (~1e-16 of the row's largest entry) over the step
    Jmax = np.max(np.abs(first(z0)), axis=1)
    for d in range(nz):
        h = 1e-3 * max(abs(z0[d]), scale[d])
        H[:, :, d] = (4.0 * central(d, 0.5 * h) - central(d, h)) / 3.0
        noise[:, :, d] = (2.2e-16 * Jmax / (0.5 * h))[:, None]
    return 0.5 * (H + H.transpose(0, 2, 1)), 0.5 * (noise + noise.transpose(0, 2, 1))


def _t(a, backend):
    return None if a is None else torch.tensor(a, device=backend)


def _row_err(got, ref):
    scale = np.max(np.abs(ref), axis=(-2, -1), keepdims=True)      # per output row
    return float(np.max(np.abs(got - ref) / np.maximum(scale, 1e-300)))


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", MODELS)
def test_exact_second_derivatives_match_the_oracle_route(backend, name):
    model, sh, X, U = _setup(name, backend)
    res = sh.hessian(_t(X, backend), _t(U, backend))
    ns, nx, nz, B = sh.n_shooting, sh.nx, sh.nx + sh.nu, X.shape[-1]
    assert res.nz == nz and tuple(res.D2.shape) == (ns, nx, nz * (nz + 1) // 2, B)
    D2 = res.full()
    assert tuple(D2.shape) == (ns, nx, nz, nz, B)
    assert torch.equal(D2, D2.transpose(2, 3))
    got = D2.cpu().numpy()
    for k, b in ((0, 0), (3, 1), (ns - 1, 2)):
        ref, noise = _oracle_second(model, sh, X, U, k, b)
        scale = np.max(np.abs(ref), axis=(1, 2), keepdims=True)     # per output row
        # 1e-10 of the row's scale (Richardson truncation) + the rounding floor of the ORACLE route (small rows: the fatigue
        # states move by ~1e-4 of an O(1) Jacobian)
        excess = np.abs(got[k, :, :, :, b] - ref) - (1e-10 * scale + 8.0 * noise)
        assert float(excess.max()) <= 0.0, (k, b, float(excess.max()), _row_err(got[k, :, :, :, b], ref))
    # structure: the Cn row is linear in everything but the intensities -> its second derivatives vanish identically
    if sh.nu <= 1:
        assert float(D2[:, 0].abs().max()) == 0.0


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name,scheme", [("ding2007_with_fatigue", "RK4"), ("ding2007_with_fatigue", "RK2"),
                                         ("ding2003_with_fatigue", "RK1"), ("hmed2018", "RK4")])
def test_exact_agrees_with_differences_of_the_exact_jacobian(backend, name, scheme):
    model, sh, X, U = _setup(name, backend, scheme=scheme, seed=9)
    exact = sh.second_derivatives(_t(X, backend), _t(U, backend))
    fd = sh.second_derivatives(_t(X, backend), _t(U, backend), method="fd")
    assert exact.shape == fd.shape
    e, f = exact.cpu().numpy(), fd.cpu().numpy()
    err = _row_err(np.moveaxis(e, -1, 0), np.moveaxis(f, -1, 0))
    assert err < 1e-7, err                                            # FD truncation + rounding, measured ~1e-9


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_multiplier_contraction_and_packed_layout(backend):
    model, sh, X, U = _setup("ding2007_with_fatigue", backend, B=5)
    ns, nx, B = sh.n_shooting, sh.nx, X.shape[-1]
    lam = torch.tensor(np.random.default_rng(1).normal(size=(ns, nx, B)), device=backend)
    res = sh.hessian(_t(X, backend), _t(U, backend), multipliers=lam)
    ref = torch.einsum("krb,krpb->kpb", lam, res.D2)
    assert float((res.H - ref).abs().max()) <= 1e-13 * float(ref.abs().max())
    only_h = sh.hessian(_t(X, backend), _t(U, backend), multipliers=lam, second=False)
    assert only_h.D2 is None and torch.equal(only_h.H, res.H)
    Hf = res.full(res.H)
    assert tuple(Hf.shape) == (ns, res.nz, res.nz, B)
    assert torch.equal(Hf, FesShooting.lagrangian_hessian(lam, res.full()).to(Hf.dtype)) or \
        float((Hf - FesShooting.lagrangian_hessian(lam, res.full())).abs().max()) <= 1e-12 * float(Hf.abs().max())
    # packed order: upper triangle, row-major
    idx = res.pair_index()
    assert idx[0, 0] == 0 and idx[0, 1] == 1 and idx[1, 1] == res.nz and idx[1, 0] == 1
    # shared parameters / shared controls give the same numbers as their expanded forms
    Ush = np.repeat(U[:, :, :1], B, axis=2)
    a = sh.hessian(_t(X, backend), _t(Ush, backend)).D2
    b = sh.hessian(_t(X, backend), _t(U[:, :, 0], backend)).D2
    assert torch.equal(a, b)


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", ["ding2003", "ding2007_with_fatigue", "hmed2018_with_fatigue"])
def test_force_factor_direction(backend, name):
    """With the force factor c of f_dot_fun (ding2003.py:309-311) as an input, z gains one direction: its row / column
    against central differences of the exact Jacobian (Jx, Ju, Jc of fes.shoot with force_factor)."""
    model, sh, X, U = _setup(name, backend)
    ns, nx, nu, B = sh.n_shooting, sh.nx, sh.nu, X.shape[-1]
    ff = np.random.default_rng(2).uniform(0.6, 1.4, (ns, B))
    res = sh.hessian(_t(X, backend), _t(U, backend), force_factor=_t(ff, backend))
    nz = nx + nu + 1
    assert res.nz == nz
    D2 = res.full().cpu().numpy()

    def jac(f):
        r = sh.evaluate(_t(X, backend), _t(U, backend), want=("x", "u"), force_factor=_t(f, backend))
        blocks = [r.Jx] + ([r.Ju] if nu else []) + [r.Jc[:, :, None]]
        return torch.cat(blocks, dim=2).cpu().numpy()                 # [ns, nx, nz, B]

    h = 1e-5
    col = (jac(ff + h) - jac(ff - h)) / (2 * h)                       # d J / d c = D2[..., :, nz-1]
    got = D2[:, :, :, nz - 1]
    scale = np.max(np.abs(col), axis=2, keepdims=True)
    assert float(np.max(np.abs(got - col) / np.maximum(scale, 1e-300))) < 1e-7
    # with c = 1 the (x, u) block is the one of the call without the factor
    ones = sh.hessian(_t(X, backend), _t(U, backend), force_factor=torch.ones((ns, B), dtype=torch.float64, device=backend)).full()
    plain = sh.hessian(_t(X, backend), _t(U, backend)).full()
    d = (ones[:, :, :nz - 1, :nz - 1] - plain).abs().max()
    assert float(d) <= 1e-13 * float(plain.abs().max())


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_argument_checks_through_the_c_abi(backend):
    model, sh, X, U = _setup("ding2007", backend)
    L = _lib.lib()
    ns, nx, B = sh.n_shooting, sh.nx, X.shape[-1]
    Xt, Ut = _t(X, backend), _t(U, backend)
    par = torch.tensor(model.param_vector(), device=backend)
    out = torch.empty((ns, nx, 6, B), dtype=torch.float64, device=backend)
    fn = L.fes_shoot_hess_f64 if backend.type == "cuda" else L.fes_shoot_hess_host_f64

    def call(dirs, d2=out, lam=None, H=None, ctrl=Ut):
        desc = fes._desc(model.kind, False, 4, sh.N, ns, 5, dirs, _lib.FES_PARAMS_SHARED | _lib.FES_WIN_SHARED, B, sh.dt, 0.0,
                         model.tauc, model.r0_km_relationship)
        buf = _lib.FesHessBuffers(fes._ptr(par), fes._ptr(Xt), fes._ptr(sh.window), fes._ptr(ctrl), fes._ptr(lam), fes._ptr(d2), fes._ptr(H))
        return fn(ctypes.byref(desc), ctypes.byref(buf), None)

    assert call(fes.FES_DIR_X0 | fes.FES_DIR_U) == 0
    assert call(fes.FES_DIR_X0) == -4                      # all of z or nothing
    assert call(fes.FES_DIR_P) == -4
    assert call(fes.FES_DIR_X0 | fes.FES_DIR_U, d2=None) == -5     # no output requested
    assert call(fes.FES_DIR_X0 | fes.FES_DIR_U, H=out) == -5       # H without multipliers
    assert call(fes.FES_DIR_X0 | fes.FES_DIR_U, ctrl=None) == -5
    assert L.fes_shoot_hess_num_dirs(_lib.FES_HMED2018, 1, 10, fes.FES_DIR_X0 | fes.FES_DIR_U | fes.FES_DIR_C) == 16
    if backend.type == "cuda":
        torch.cuda.synchronize()


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_edge_shapes_single_unit_long_and_per_scenario_windows(backend):
    """One scenario / one step / one-slot window; a window left-padded to 300 slots (the truncation sweep's layout) gives the
    numbers of the short one bit for bit (pads add exact zeros); a per-scenario window equals the shared one."""
    kind, fat = fes.FES_DING2003, True
    par = torch.tensor(cf.ModelMaker.create_model("ding2003_with_fatigue", stim_time=[0.0], sum_stim_truncation=1).param_vector(), device=backend)
    X = torch.tensor([[[0.3], [40.0], [2800.0], [0.055], [0.11]]] * 2, dtype=torch.float64, device=backend)     # [2, 5, 1]
    win1 = torch.tensor([[0.0]], dtype=torch.float64, device=backend)
    one = fes.shoot_hessian(kind, fat, par, X, win1, 0.05, n_steps=1)
    assert tuple(one.D2.shape) == (1, 5, 15, 1) and bool(torch.isfinite(one.D2).all())
    ten = fes.shoot_hessian(kind, fat, par, X, win1, 0.05, n_steps=10)
    assert float((one.D2 - ten.D2).abs().max()) > 0.0               # the discrete map, not the continuous one
    short = torch.tensor([[-0.1, -0.05, 0.0]], dtype=torch.float64, device=backend)
    long = torch.cat([torch.full((1, 297), -1.0e7, dtype=torch.float64, device=backend), short], dim=1)
    a = fes.shoot_hessian(kind, fat, par, X, short, 0.05, n_steps=4).D2
    b = fes.shoot_hessian(kind, fat, par, X, long, 0.05, n_steps=4).D2
    assert torch.equal(a, b)
    B = 4
    XB = X.repeat(1, 1, B) * torch.linspace(0.9, 1.1, B, dtype=torch.float64, device=backend)
    shared = fes.shoot_hessian(kind, fat, par, XB, short, 0.05, n_steps=4).D2
    per = fes.shoot_hessian(kind, fat, par[:, None].repeat(1, B), XB, short[:, :, None].repeat(1, 1, B), 0.05, n_steps=4).D2
    assert torch.equal(shared, per)
    # pads only: no pulse has happened, the Cn row is pure decay and the blocks stay finite
    pads = torch.full((1, 3), -1.0e7, dtype=torch.float64, device=backend)
    assert bool(torch.isfinite(fes.shoot_hessian(kind, fat, par, X, pads, 0.05, n_steps=4).D2).all())
