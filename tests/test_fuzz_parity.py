"""Seeded random sweep over the shape space of the kernels (model, scheme, truncation N, intervals, sub-steps, batch,
pulse frequency, shared / per-trajectory inputs, direction group) against the C oracle: forward map, chained IVP
sensitivities, shooting Jacobian blocks.  Complements the fixed cases of test_parity_*: every case exercises the stage
table (eps, psi), the shared Cn modes, the reduced fatigue tangents and -- for Hmed2018 -- the tanh cache."""
import numpy as np
import pytest
import torch

from cocofest_b200 import fes
from oracle import fes_oracle as fo
from oracle import fes_oracle_c as foc
from tests.conftest import BACKENDS
from tests.problems import MODELS, P_COLS, PF_COLS, oracle_win, random_problem, rel_err, to_lib

RTOL64 = 1e-9
NAMES = list(MODELS)


def _draw(case):
    rng = np.random.default_rng(1000 + case)
    name = NAMES[case % len(NAMES)]
    freq = float(rng.choice([10.0, 20.0, 33.0, 50.0]))
    # h <= 0.01 s keeps h / tauc <= 0.91 (tauc = 0.011 for Ding2007): inside the stability region of all three schemes, so
    # rounding differences are not amplified (an unstable RK1 run diverges to 1e11 and no tolerance means anything)
    S = int(np.ceil(1.0 / freq / 0.01)) + int(rng.integers(0, 4))
    return dict(name=name, scheme=int(rng.choice([1, 2, 4])), N=int(rng.integers(1, 13)), ns=int(rng.integers(1, 12)),
                S=S, B=int(rng.integers(1, 75)), freq=freq,
                shared_params=bool(rng.integers(0, 2)), shared_ctrl=bool(rng.integers(0, 2)),
                per_traj_win=bool(rng.integers(0, 2)), seed=int(rng.integers(0, 10_000)))


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("case", range(18))
def test_random_ivp_case(backend, case):
    c = _draw(case)
    pb = random_problem(c["name"], B=c["B"], N=c["N"], ns=c["ns"], freq=c["freq"], seed=c["seed"],
                        shared_params=c["shared_params"], shared_ctrl=c["shared_ctrl"], per_traj_win=c["per_traj_win"])
    params, x0, win, ctrl = to_lib(pb, backend)
    use_p = case % 3 == 2
    with_pf = use_p and pb["fat"] and case % 2 == 0
    dirs = (fes.FES_DIR_P | (fes.FES_DIR_PF if with_pf else 0)) if use_p else fes.FES_DIR_X0
    res = fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], n_steps=c["S"], ctrl=ctrl, n_shooting=pb["ns"],
                  scheme=c["scheme"], dirs=dirs, trace="nodes")
    pdir = (P_COLS[pb["kind"]] + (PF_COLS if with_pf else [])) if use_p else ()
    ref = foc.ivp(pb["kind"], pb["fat"], c["scheme"], pb["N"], pb["ns"], c["S"], pb["dt"], pb["tauc"], fo.R0_KM_RELATIONSHIP,
                  pb["params"], pb["x0"], oracle_win(pb), pb["u"], seed_x0=not use_p, pdir=pdir, trace_B=pb["B"],
                  trace_every=c["S"])
    assert rel_err(res.x_final.cpu().numpy().T, ref["final"]) < RTOL64, c
    assert rel_err(np.transpose(res.trace.cpu().numpy(), (2, 0, 1)), ref["trace"]) < RTOL64, c
    got = np.transpose(res.sens_final.cpu().numpy(), (2, 0, 1))
    for col in range(got.shape[2]):
        assert rel_err(got[:, :, col], ref["sens_final"][:, :, col], floor=1e-9) < 10 * RTOL64, (c, res.columns[col])


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("case", range(12))
def test_random_shoot_case(backend, case):
    c = _draw(100 + case)
    pb = random_problem(c["name"], B=c["B"], N=c["N"], ns=c["ns"], freq=c["freq"], seed=c["seed"],
                        shared_params=c["shared_params"], shared_ctrl=c["shared_ctrl"], per_traj_win=c["per_traj_win"])
    rng = np.random.default_rng(c["seed"])
    B, ns, nx = pb["B"], pb["ns"], pb["nx"]
    X = np.zeros((B, ns + 1, nx))
    X[:, :, 0] = rng.uniform(0.02, 1.0, (B, ns + 1))
    X[:, :, 1] = rng.uniform(1.0, 250.0, (B, ns + 1))
    if pb["fat"]:
        pp = np.broadcast_to(pb["params"], (B, fo.NP_FULL))
        X[:, :, 2] = pp[:, None, fo.P_A_REST] * rng.uniform(0.6, 1.0, (B, ns + 1))
        X[:, :, 3] = pp[:, None, fo.P_TAU1_REST] * rng.uniform(1.0, 1.5, (B, ns + 1))
        X[:, :, 4] = pp[:, None, fo.P_KM_REST] * rng.uniform(1.0, 1.5, (B, ns + 1))
    params, _, win, ctrl = to_lib(pb, backend)
    Xl = torch.tensor(np.ascontiguousarray(np.transpose(X, (1, 2, 0))), device=backend)
    want = ["x", "u", "p", "w"] + (["pf"] if pb["fat"] else [])
    res = fes.shoot(pb["kind"], pb["fat"], params, Xl, win, pb["dt"], n_steps=c["S"], ctrl=ctrl, scheme=c["scheme"], want=want)
    pdir = P_COLS[pb["kind"]] + (PF_COLS if pb["fat"] else [])
    ref = foc.shoot(pb["kind"], pb["fat"], c["scheme"], pb["N"], ns, c["S"], pb["dt"], 0.0, pb["tauc"], fo.R0_KM_RELATIONSHIP,
                    pb["params"], X, oracle_win(pb), pb["u"], pdir=pdir, want_x=True, want_u=pb["nu"] > 0, want_w=True)
    bl = lambda t: np.transpose(t.cpu().numpy(), (3, 0, 1, 2))
    scale = np.max(np.abs(X[:, 1:]))
    assert np.max(np.abs(np.transpose(res.G.cpu().numpy(), (2, 0, 1)) - ref["G"])) < RTOL64 * scale, c
    assert rel_err(bl(res.Jx), ref["Jx"], floor=1e-9) < 10 * RTOL64, c
    if pb["nu"]:
        assert rel_err(bl(res.Ju), ref["Ju"], floor=1e-9) < 10 * RTOL64, c
    Jp = bl(res.Jp)
    for col in range(Jp.shape[3]):
        assert rel_err(Jp[..., col], ref["Jp"][..., col], floor=1e-9) < 10 * RTOL64, (c, res.p_columns[col])
    assert rel_err(bl(res.Jw), ref["Jw"], floor=1e-9) < 10 * RTOL64, c
