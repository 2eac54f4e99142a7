"""Force-tracking cost / gradient / Gauss-Newton matrix and the on-device Levenberg-Marquardt identification
(fes_identify_*) against the oracle (oracle/fes_oracle_ident.py on top of the pinned C forward map).

Every test runs twice: `emul` (host build of the kernel source, CPU CI) and `cuda` (the product, -m gpu).
Tolerance: 1e-9 relative in fp64 for the evaluation; the identified parameters are compared with the numpy
restatement of the same iteration and with the parameters that generated the data."""
import numpy as np
import pytest
import torch

from cocofest_b200 import fes
from cocofest_b200._lib import FesLibraryError
from oracle import fes_oracle as fo
from oracle import fes_oracle_ident as foi
from tests.conftest import BACKENDS
from tests.problems import entry_err, rel_err

KINDS = {"ding2003": fo.DING2003, "ding2007": fo.DING2007, "hmed2018": fo.HMED2018}
# default boxes / initial guesses of the reference identification classes
BOXES = {
    fo.DING2003: dict(lo={"a_rest": 1, "km_rest": 0.001, "tau1_rest": 0.0001, "tau2": 0.0001},
                      hi={"a_rest": 10000, "km_rest": 1, "tau1_rest": 1, "tau2": 1},
                      guess={"a_rest": 1000, "km_rest": 0.5, "tau1_rest": 0.5, "tau2": 0.5}),
}


def _protocol(kind, seed=0, N=10, final_time=2.0):
    """examples/identification/*_model_identification.py: 10 pulses at 10 Hz, observed for 2 s."""
    rng = np.random.default_rng(seed)
    stim = list(np.linspace(0, 1, 11)[:-1])
    ns = fes.prepare_n_shooting(stim, final_time)
    win, idx = fes.stim_windows(stim, ns, final_time, N)
    u = None
    if kind == fo.DING2007:
        u = fo.ivp_controls_pulse_width(list(rng.uniform(2e-4, 6e-4, len(stim))), idx, ns)[None]
    if kind == fo.HMED2018:
        u = fo.ivp_controls_pulse_intensity(list(rng.integers(30, 130, len(stim)).astype(float)), idx, ns, N)[None]
    return dict(stim=stim, ns=ns, dt=final_time / ns, win=win, u=u, N=N, tauc=fo.default_tauc(kind))


def _ctrl_tensor(pr, device):
    return None if pr["u"] is None else torch.tensor(np.ascontiguousarray(pr["u"][0]), device=device)


def _vec(d, fill):
    v = np.full(fo.NP_FULL, fill)
    for k, val in d.items():
        v[fes.PARAM_INDEX[k]] = val
    return v


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", list(KINDS))
@pytest.mark.parametrize("shared_target", [True, False])
def test_cost_gradient_hessian_match_oracle(backend, name, shared_target):
    kind = KINDS[name]
    pr = _protocol(kind, seed=1)
    rng = np.random.default_rng(2)
    B, ns = 9, pr["ns"]
    params = fo.default_params(kind)[None] * rng.uniform(0.8, 1.25, (B, fo.NP_FULL))
    F_true, _ = foi.node_forces(kind, fo.default_params(kind)[None], pr["win"][None], pr["u"], ns, 10, pr["dt"], pr["tauc"])
    target = F_true if shared_target else F_true * rng.uniform(0.9, 1.1, (B, ns + 1))
    weight = rng.uniform(0.5, 1.5, ns + 1) * pr["dt"]
    x0 = np.stack([rng.uniform(0, 0.3, B), rng.uniform(0, 40, B)], axis=1)
    c, g, H = foi.cost_grad_hess(kind, params, pr["win"][None], pr["u"], target, ns, 10, pr["dt"], pr["tauc"], weight=weight, x0=x0)
    res = fes.identify(kind, torch.tensor(params.T.copy(), device=backend), torch.tensor(pr["win"], device=backend), pr["dt"],
                       torch.tensor(target[0] if shared_target else target.T.copy(), device=backend), ctrl=_ctrl_tensor(pr, backend),
                       weight=torch.tensor(weight, device=backend), x0=torch.tensor(x0.T.copy(), device=backend), max_iter=0)
    assert rel_err(res.cost.cpu().numpy(), c) < 1e-9
    assert entry_err(res.grad.cpu().numpy().T, g, keep=(1,)) < 1e-9                                   # per parameter
    assert entry_err(np.transpose(res.hessian().cpu().numpy(), (2, 0, 1)), H, keep=(1, 2)) < 1e-9      # per (d, e) entry
    np.testing.assert_array_equal(res.params.cpu().numpy(), params.T)  # max_iter = 0 leaves the candidates alone
    assert (res.evaluations.cpu().numpy() == 1).all()


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", list(KINDS))
@pytest.mark.parametrize("shared_target", [True, False])
def test_cost_gradient_hessian_fp32_mode(backend, name, shared_target):
    """fes_identify_f32 against the fp64 oracle at north_star's 1e-4: cost per candidate, every gradient / Gauss-Newton
    entry on its own scale."""
    kind = KINDS[name]
    pr = _protocol(kind, seed=1)
    rng = np.random.default_rng(2)
    B, ns = 9, pr["ns"]
    params = fo.default_params(kind)[None] * rng.uniform(0.8, 1.25, (B, fo.NP_FULL))
    F_true, _ = foi.node_forces(kind, fo.default_params(kind)[None], pr["win"][None], pr["u"], ns, 10, pr["dt"], pr["tauc"])
    target = F_true if shared_target else F_true * rng.uniform(0.9, 1.1, (B, ns + 1))
    weight = rng.uniform(0.5, 1.5, ns + 1) * pr["dt"]
    x0 = np.stack([rng.uniform(0, 0.3, B), rng.uniform(0, 40, B)], axis=1)
    c, g, H = foi.cost_grad_hess(kind, params, pr["win"][None], pr["u"], target, ns, 10, pr["dt"], pr["tauc"], weight=weight, x0=x0)
    f32 = torch.float32
    u = _ctrl_tensor(pr, backend)
    res = fes.identify(kind, torch.tensor(params.T.copy(), dtype=f32, device=backend), torch.tensor(pr["win"], device=backend),
                       pr["dt"], torch.tensor(target[0] if shared_target else target.T.copy(), dtype=f32, device=backend),
                       ctrl=None if u is None else u.to(f32), weight=torch.tensor(weight, dtype=f32, device=backend),
                       x0=torch.tensor(x0.T.copy(), dtype=f32, device=backend), max_iter=0, dtype=f32)
    assert res.cost.dtype == f32
    assert rel_err(res.cost.cpu().numpy(), c) < 1e-4
    assert entry_err(res.grad.cpu().numpy().T, g, keep=(1,)) < 1e-4
    assert entry_err(np.transpose(res.hessian().cpu().numpy(), (2, 0, 1)), H, keep=(1, 2)) < 1e-4


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_gradient_is_the_derivative_of_the_cost(backend):
    """Central differences of the library's own cost (property, independent of the oracle)."""
    kind = fo.HMED2018
    pr = _protocol(kind, seed=3)
    p0 = fo.default_params(kind) * 1.1
    cols = foi.P_COLS[kind]
    rel = 1e-6
    pts = [p0.copy()]
    for j in cols:
        for sgn in (+1, -1):
            q = p0.copy()
            q[j] *= 1 + sgn * rel
            pts.append(q)
    target, _ = foi.node_forces(kind, fo.default_params(kind)[None], pr["win"][None], pr["u"], pr["ns"], 10, pr["dt"], pr["tauc"])
    res = fes.identify(kind, torch.tensor(np.array(pts).T.copy(), device=backend), torch.tensor(pr["win"], device=backend),
                       pr["dt"], torch.tensor(target[0], device=backend), ctrl=_ctrl_tensor(pr, backend), max_iter=0)
    cost = res.cost.cpu().numpy()
    g = res.grad.cpu().numpy()[:, 0]
    for d, j in enumerate(cols):
        fd = (cost[1 + 2 * d] - cost[2 + 2 * d]) / (2 * rel * p0[j])
        assert abs(fd - g[d]) <= 1e-5 * max(abs(g[d]), 1e-12), (d, fd, g[d])


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_ding2003_recovers_generating_parameters_from_reference_guess(backend):
    """examples/identification/ding2003_model_identification.py without noise: data generated at the default
    parameters, started from the reference's default initial guess inside its default box."""
    kind = fo.DING2003
    pr = _protocol(kind)
    truth = fo.default_params(kind)
    target, _ = foi.node_forces(kind, truth[None], pr["win"][None], None, pr["ns"], 10, pr["dt"], pr["tauc"])
    box = BOXES[kind]
    p0 = truth.copy()
    for k, v in box["guess"].items():
        p0[fes.PARAM_INDEX[k]] = v
    rng = np.random.default_rng(5)
    starts = np.stack([p0] + [p0 * np.where(np.arange(fo.NP_FULL) < 4, rng.uniform(0.5, 1.5, fo.NP_FULL), 1.0) for _ in range(5)])
    res = fes.identify(kind, torch.tensor(starts.T.copy(), device=backend), torch.tensor(pr["win"], device=backend), pr["dt"],
                       torch.tensor(target[0], device=backend), lower=box["lo"], upper=box["hi"], max_iter=200)
    b = res.best()
    got = res.params.cpu().numpy()[:, b]
    assert float(res.cost[b]) < 1e-12 * float(np.sum(target ** 2) * pr["dt"])
    np.testing.assert_allclose(got[:4], truth[:4], rtol=1e-6)
    np.testing.assert_array_equal(got[4:], truth[4:])
    # same iteration restated in numpy on the oracle's forward map, candidate 0
    ref = foi.levenberg_marquardt(kind, starts[0], pr["win"][None], None, target, pr["ns"], 10, pr["dt"], pr["tauc"],
                                  lower=_vec(box["lo"], -np.inf), upper=_vec(box["hi"], np.inf), max_iter=200)
    np.testing.assert_allclose(res.params.cpu().numpy()[:4, 0], ref["params"][:4], rtol=1e-6)


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", ["ding2007", "hmed2018"])
def test_short_iteration_follows_numpy_restatement(backend, name):
    """A few accepted / rejected steps from a nearby start: iterate-level agreement with the numpy LM."""
    kind = KINDS[name]
    pr = _protocol(kind, seed=7)
    truth = fo.default_params(kind)
    target, _ = foi.node_forces(kind, truth[None], pr["win"][None], pr["u"], pr["ns"], 10, pr["dt"], pr["tauc"])
    rng = np.random.default_rng(8)
    B = 4
    starts = truth[None] * rng.uniform(0.9, 1.1, (B, fo.NP_FULL))
    ident = ["a_scale", "km_rest", "tau1_rest", "pd0"] if kind == fo.DING2007 else ["a_rest", "tau1_rest", "ar", "Is"]
    res = fes.identify(kind, torch.tensor(starts.T.copy(), device=backend), torch.tensor(pr["win"], device=backend), pr["dt"],
                       torch.tensor(target[0], device=backend), ctrl=_ctrl_tensor(pr, backend), identify=ident, max_iter=4)
    canon = [fes.PARAM_INDEX[k] for k in ident]
    got = res.params.cpu().numpy()
    for b in range(B):
        ref = foi.levenberg_marquardt(kind, starts[b], pr["win"][None], pr["u"], target, pr["ns"], 10, pr["dt"], pr["tauc"],
                                      identify=canon, max_iter=4)
        np.testing.assert_allclose(got[:, b], ref["params"], rtol=1e-7)
        assert int(res.evaluations[b]) == ref["evaluations"]
        assert abs(float(res.cost[b]) - ref["cost"]) <= 1e-7 * max(ref["cost"], 1e-30)
        fixed = [j for j in range(fo.NP_FULL) if j not in canon]
        np.testing.assert_array_equal(got[fixed, b], starts[b, fixed])


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_per_candidate_targets_with_shared_start(backend):
    """B data sets (one per subject), one shared starting point: params [14], target [ns+1, B]."""
    kind = fo.DING2003
    pr = _protocol(kind)
    rng = np.random.default_rng(11)
    B = 5
    truths = fo.default_params(kind)[None] * np.where(np.arange(fo.NP_FULL) < 4, rng.uniform(0.8, 1.2, (B, fo.NP_FULL)), 1.0)
    target, _ = foi.node_forces(kind, truths, pr["win"][None], None, pr["ns"], 10, pr["dt"], pr["tauc"])
    res = fes.identify(kind, torch.tensor(fo.default_params(kind), device=backend), torch.tensor(pr["win"], device=backend),
                       pr["dt"], torch.tensor(target.T.copy(), device=backend), max_iter=100)
    np.testing.assert_allclose(res.params.cpu().numpy()[:4].T, truths[:, :4], rtol=1e-6)


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_box_is_respected(backend):
    kind = fo.DING2003
    pr = _protocol(kind)
    truth = fo.default_params(kind)
    target, _ = foi.node_forces(kind, truth[None], pr["win"][None], None, pr["ns"], 10, pr["dt"], pr["tauc"])
    lo, hi = {"a_rest": 3100.0, "tau2": 0.01}, {"a_rest": 5000.0, "tau2": 0.07}
    res = fes.identify(kind, torch.tensor((truth * 1.2)[:, None].copy(), device=backend), torch.tensor(pr["win"], device=backend),
                       pr["dt"], torch.tensor(target[0], device=backend), lower=lo, upper=hi, max_iter=400)
    got = res.params.cpu().numpy()[:, 0]
    assert 3100.0 <= got[0] <= 5000.0 and 0.01 <= got[3] <= 0.07
    assert got[0] == pytest.approx(3100.0)  # the unconstrained optimum (3009) lies below the box
    # first-order optimality on the free parameters, outward gradient on the held one
    g = res.grad.cpu().numpy()[:, 0]
    assert g[0] > 0.0
    ref = foi.levenberg_marquardt(kind, truth * 1.2, pr["win"][None], None, target, pr["ns"], 10, pr["dt"], pr["tauc"],
                                  lower=_vec(lo, -np.inf), upper=_vec(hi, np.inf), max_iter=400)
    # non-zero residual at the constrained optimum: Gauss-Newton converges linearly and the ftol stop is
    # rounding-sensitive, so the two iterations agree to the stopping tolerance, not to 1e-9
    np.testing.assert_allclose(got[:4], ref["params"][:4], rtol=1e-4)
    assert abs(float(res.cost[0]) - ref["cost"]) <= 1e-8 * ref["cost"]


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_fixed_parameter_outside_a_full_tensor_box_is_left_alone(backend):
    """[14] tensor bounds cover every parameter; one that is NOT identified and starts outside its box must keep its
    starting value (it is neither projected at the start nor clamped by a trial step) and must not count as a step."""
    kind = fo.DING2003
    pr = _protocol(kind)
    truth = fo.default_params(kind)
    target, _ = foi.node_forces(kind, truth[None], pr["win"][None], None, pr["ns"], 10, pr["dt"], pr["tauc"])
    lo, hi = truth * 0.5, truth * 1.5
    lo[fo.P_TAU2], hi[fo.P_TAU2] = 0.02, 0.04                # tau2 = 0.06 is fixed and lies above its box
    lo, hi = np.minimum(lo, hi), np.maximum(lo, hi)          # alpha_a is negative
    start = truth * np.where(np.arange(fo.NP_FULL) < 3, 1.15, 1.0)
    ident = ["a_rest", "km_rest", "tau1_rest"]
    res = fes.identify(kind, torch.tensor(start[:, None].copy(), device=backend), torch.tensor(pr["win"], device=backend), pr["dt"],
                       torch.tensor(target[0], device=backend), identify=ident, lower=torch.tensor(lo), upper=torch.tensor(hi),
                       max_iter=100)
    got = res.params.cpu().numpy()[:, 0]
    assert got[fo.P_TAU2] == truth[fo.P_TAU2]                # untouched
    np.testing.assert_allclose(got[:3], truth[:3], rtol=1e-6)   # and the identified ones reach the generating values
    ref = foi.levenberg_marquardt(kind, start, pr["win"][None], None, target, pr["ns"], 10, pr["dt"], pr["tauc"],
                                  identify=[fes.PARAM_INDEX[k] for k in ident], lower=lo, upper=hi, max_iter=100)
    # (near a zero-residual optimum the stopping tests are rounding-sensitive: the iterates agree, the evaluation counts need not)
    assert ref["params"][fo.P_TAU2] == truth[fo.P_TAU2]
    np.testing.assert_allclose(got[:3], ref["params"][:3], rtol=1e-6)


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_argument_errors(backend):
    kind = fo.DING2003
    pr = _protocol(kind)
    p = torch.tensor(fo.default_params(kind)[:, None].copy(), device=backend)
    win = torch.tensor(pr["win"], device=backend)
    tgt = torch.zeros(pr["ns"] + 1, dtype=torch.float64, device=backend)
    with pytest.raises(ValueError):
        fes.identify(kind, p, win, pr["dt"], tgt, identify=["pd0"])  # not a ding2003 parameter
    with pytest.raises(ValueError):
        fes.identify(kind, p[:, 0], win, pr["dt"], tgt)  # no batch axis anywhere
    with pytest.raises(FesLibraryError):
        fes.identify(kind, p, win, pr["dt"], tgt, lambda0=0.0)
