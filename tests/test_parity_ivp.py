"""Parity of the kernel arithmetic against the oracle: batched IVP (forward map + chained sensitivities).

Every test runs twice: `emul` (host build of the kernel source, CPU CI) and `cuda` (the product, -m gpu).
Tolerance: 1e-9 relative in fp64 (BASELINE.json north_star), 1e-4 in fp32."""
import json
from pathlib import Path

import numpy as np
import pytest
import torch

from cocofest_b200 import fes
from oracle import fes_oracle as fo
from oracle import fes_oracle_c as foc
from tests.conftest import BACKENDS
from tests.problems import MODELS, P_COLS, PF_COLS, entry_err, oracle_win, random_problem, rel_err, to_lib

GOLD = json.loads((Path(__file__).parent / "golden" / "ivp_golden.json").read_text())
RTOL64 = 1e-9
RTOL32 = 1e-4


def _golden_inputs(name, device, pulse_mode="single", stim=(0, 0.1, 0.2)):
    kind, fat = MODELS[name]
    st = fes.pulse_mode_times(list(stim), pulse_mode)
    ns = fes.prepare_n_shooting(st, 0.3)
    win, idx = fes.stim_windows(st, ns, 0.3, 3)
    p = fes.default_params(kind)
    ctrl = None
    if kind == fo.DING2007:
        ctrl = torch.tensor(fo.ivp_controls_pulse_width(GOLD["pulse_width"], idx, ns), device=device)
    if kind == fo.HMED2018:
        ctrl = torch.tensor(fo.ivp_controls_pulse_intensity(GOLD["pulse_intensity"], idx, ns, 3), device=device)
    x0 = torch.tensor(fo.rest_state(kind, fat, p), device=device)[:, None]
    return kind, fat, torch.tensor(p, device=device), x0, torch.tensor(win, device=device), ctrl, ns


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", list(MODELS))
def test_reference_golden_force_trajectories(backend, name):
    """The six 31-point force vectors of reference tests/shard1/test_ivp.py:28-124."""
    kind, fat, p, x0, win, ctrl, ns = _golden_inputs(name, backend)
    res = fes.ivp(kind, fat, p, x0, win, 0.3 / ns, n_steps=10, ctrl=ctrl, n_shooting=ns, trace="steps")
    F = res.trace[:, 1, 0].cpu().numpy()
    np.testing.assert_almost_equal(F, np.array(GOLD["F"][name]), decimal=GOLD["decimal"])


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_reference_golden_pulse_modes(backend):
    """Doublet and (state-leaked) triplet slices of test_ivp.py:127-169."""
    name = "ding2003_with_fatigue"
    kind, fat, p, x0, win, ctrl, ns = _golden_inputs(name, backend, "doublet")
    assert ns == 60
    res = fes.ivp(kind, fat, p, x0, win, 0.3 / ns, n_steps=10, n_shooting=ns, trace="steps")
    np.testing.assert_almost_equal(res.trace[150:180, 1, 0].cpu().numpy(),
                                   np.array(GOLD["F"]["pulse_mode_doublet_150_180"]), decimal=7)
    leaked = fes.pulse_mode_times(fes.pulse_mode_times([0, 0.1, 0.2], "doublet"), "triplet")
    kind, fat, p, x0, win, ctrl, ns = _golden_inputs(name, backend, "single", stim=leaked)
    res = fes.ivp(kind, fat, p, x0, win, 0.3 / ns, n_steps=10, n_shooting=ns, trace="steps")
    np.testing.assert_almost_equal(res.trace[350:380, 1, 0].cpu().numpy(),
                                   np.array(GOLD["F"]["pulse_mode_triplet_350_380"]), decimal=7)


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", list(MODELS))
@pytest.mark.parametrize("scheme", [1, 2, 4])
def test_ivp_forward_matches_oracle(backend, name, scheme):
    pb = random_problem(name, B=37, seed=11, shared_ctrl=False)
    S = 5
    params, x0, win, ctrl = to_lib(pb, backend)
    res = fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], n_steps=S, ctrl=ctrl, n_shooting=pb["ns"],
                  scheme=scheme, trace="steps", trace_batch=5)
    ref = foc.ivp(pb["kind"], pb["fat"], scheme, pb["N"], pb["ns"], S, pb["dt"], pb["tauc"], fo.R0_KM_RELATIONSHIP,
                  pb["params"], pb["x0"], oracle_win(pb), pb["u"], trace_B=5, trace_every=1)
    assert rel_err(res.x_final.cpu().numpy().T, ref["final"]) < RTOL64
    got = np.transpose(res.trace.cpu().numpy(), (2, 0, 1))  # [tB, npts, nx]
    assert rel_err(got, ref["trace"]) < RTOL64
    # node traces are the sub-step trace sampled every S points
    res_n = fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], n_steps=S, ctrl=ctrl, n_shooting=pb["ns"],
                    scheme=scheme, trace="nodes")
    assert torch.equal(res_n.trace[:, :, :5], res.trace[::S])
    assert torch.equal(res_n.x_final, res.x_final)


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", list(MODELS))
def test_ivp_per_trajectory_windows_and_shared_params(backend, name):
    pb = random_problem(name, B=19, seed=5, shared_params=True, per_traj_win=True)
    params, x0, win, ctrl = to_lib(pb, backend)
    res = fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], n_steps=4, ctrl=ctrl, n_shooting=pb["ns"])
    ref = foc.ivp(pb["kind"], pb["fat"], 4, pb["N"], pb["ns"], 4, pb["dt"], pb["tauc"], fo.R0_KM_RELATIONSHIP,
                  pb["params"], pb["x0"], oracle_win(pb), pb["u"])
    assert rel_err(res.x_final.cpu().numpy().T, ref["final"]) < RTOL64


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", list(MODELS))
def test_ivp_sensitivities_wrt_x0_match_oracle(backend, name):
    pb = random_problem(name, B=21, seed=3)
    S = 6
    params, x0, win, ctrl = to_lib(pb, backend)
    res = fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], n_steps=S, ctrl=ctrl, n_shooting=pb["ns"],
                  dirs=fes.FES_DIR_X0, sens_trace=True, trace="nodes", trace_batch=4)
    ref = foc.ivp(pb["kind"], pb["fat"], 4, pb["N"], pb["ns"], S, pb["dt"], pb["tauc"], fo.R0_KM_RELATIONSHIP,
                  pb["params"], pb["x0"], oracle_win(pb), pb["u"], seed_x0=True, trace_B=4, sens_trace=True)
    assert rel_err(res.x_final.cpu().numpy().T, ref["final"]) < RTOL64
    got = np.transpose(res.sens_final.cpu().numpy(), (2, 0, 1))  # [B, nx, nd]
    # every (state row, direction column) entry on its own scale: the columns differ by orders of magnitude
    assert entry_err(got, ref["sens_final"], keep=(1, 2)) < RTOL64
    got_t = np.transpose(res.sens_trace.cpu().numpy(), (3, 0, 1, 2))  # [tB, ns+1, nx, nd]
    assert entry_err(got_t[:, 1:], ref["sens_trace"][:, 1:], keep=(1, 2, 3)) < RTOL64   # per node as well
    assert np.array_equal(got_t[:, 0], ref["sens_trace"][:, 0])                          # the seed is the identity
    assert len(res.columns) == pb["nx"]


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", list(MODELS))
@pytest.mark.parametrize("with_pf", [False, True])
def test_ivp_sensitivities_wrt_parameters_match_oracle(backend, name, with_pf):
    pb = random_problem(name, B=13, seed=9, shared_ctrl=False)
    if with_pf and not pb["fat"]:
        pytest.skip("fatigue parameters only exist on fatigue models")
    S = 5
    params, x0, win, ctrl = to_lib(pb, backend)
    dirs = fes.FES_DIR_P | (fes.FES_DIR_PF if with_pf else 0)
    res = fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], n_steps=S, ctrl=ctrl, n_shooting=pb["ns"], dirs=dirs)
    pdir = P_COLS[pb["kind"]] + (PF_COLS if with_pf else [])
    ref = foc.ivp(pb["kind"], pb["fat"], 4, pb["N"], pb["ns"], S, pb["dt"], pb["tauc"], fo.R0_KM_RELATIONSHIP,
                  pb["params"], pb["x0"], oracle_win(pb), pb["u"], pdir=pdir)
    got = np.transpose(res.sens_final.cpu().numpy(), (2, 0, 1))
    assert got.shape == ref["sens_final"].shape
    # column-wise relative error: columns differ by many orders of magnitude (d/dpd0 ~ 1e6, d/dalpha ~ 1e2)
    for c in range(got.shape[2]):
        assert entry_err(got[:, :, c], ref["sens_final"][:, :, c], keep=(1,)) < RTOL64, (name, res.columns[c])


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", list(MODELS))
@pytest.mark.parametrize("group", ["x0", "p", "p|pf"])
def test_ivp_fp32_mode(backend, name, group):
    """north_star's fp32 mode: states AND sensitivities within 1e-4 of the fp64 oracle, every state on its own scale and
    every (row, column) sensitivity entry on its own scale, for the six models and every IVP direction set."""
    pb = random_problem(name, B=16, seed=2, shared_ctrl=False)
    if group == "p|pf" and not pb["fat"]:
        pytest.skip("fatigue parameters only exist on fatigue models")
    dirs = {"x0": fes.FES_DIR_X0, "p": fes.FES_DIR_P, "p|pf": fes.FES_DIR_P | fes.FES_DIR_PF}[group]
    params, x0, win, ctrl = to_lib(pb, backend, dtype=torch.float32)
    res = fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], n_steps=10, ctrl=ctrl, n_shooting=pb["ns"],
                  dirs=dirs, trace="nodes", sens_trace=True)
    assert res.x_final.dtype == torch.float32 and res.sens_final.dtype == torch.float32
    kw = dict(seed_x0=True) if group == "x0" else dict(pdir=P_COLS[pb["kind"]] + (PF_COLS if group == "p|pf" else []))
    ref = foc.ivp(pb["kind"], pb["fat"], 4, pb["N"], pb["ns"], 10, pb["dt"], pb["tauc"], fo.R0_KM_RELATIONSHIP,
                  pb["params"], pb["x0"], oracle_win(pb), pb["u"], trace_B=pb["B"], sens_trace=True, **kw)
    assert entry_err(res.x_final.cpu().numpy().T, ref["final"], keep=(1,)) < RTOL32
    assert entry_err(np.transpose(res.trace.cpu().numpy(), (2, 0, 1)), ref["trace"], keep=(2,)) < RTOL32
    got = np.transpose(res.sens_final.cpu().numpy(), (2, 0, 1))
    assert entry_err(got, ref["sens_final"], keep=(1, 2)) < RTOL32
    got_t = np.transpose(res.sens_trace.cpu().numpy(), (3, 0, 1, 2))
    assert entry_err(got_t[:, 1:], ref["sens_trace"][:, 1:], keep=(2, 3)) < RTOL32


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_ivp_long_fatigue_protocol_matches_oracle(backend):
    """BASELINE config 2 protocol (33 Hz, 1 s on / 1 s off) for 20 s: 660 intervals x 10 steps, N = 10."""
    kind, fat = MODELS["ding2003_with_fatigue"]
    T = 20
    stim = [s + k / 33 for s in range(0, T, 2) for k in range(33)]
    ns = 33 * T
    win, _ = fes.stim_windows(stim, ns, T, 10)
    B = 8
    rng = np.random.default_rng(0)
    params = fes.default_params(kind)[None] * np.concatenate([rng.uniform(0.8, 1.2, (B, 8)), np.ones((B, 6))], axis=1)
    x0 = np.zeros((B, 5))
    x0[:, 2], x0[:, 3], x0[:, 4] = params[:, 0], params[:, 2], params[:, 1]
    res = fes.ivp(kind, fat, torch.tensor(params.T.copy(), device=backend), torch.tensor(x0.T.copy(), device=backend),
                  torch.tensor(win, device=backend), T / ns, n_steps=10, n_shooting=ns, trace="nodes", dirs=fes.FES_DIR_X0,
                  sens_trace=True)
    ref = foc.ivp(kind, fat, 4, 10, ns, 10, T / ns, 0.020, 1.04, params, x0, win[None], None, seed_x0=True, trace_B=B,
                  sens_trace=True)
    got = np.transpose(res.trace.cpu().numpy(), (2, 0, 1))
    assert rel_err(got[:, :, 1], ref["trace"][:, :, 1], floor=1e-6) < RTOL64   # F at every node
    assert entry_err(res.x_final.cpu().numpy().T, ref["final"], keep=(1,)) < RTOL64
    gs = np.transpose(res.sens_final.cpu().numpy(), (2, 0, 1))
    # every (row, column) entry on its own scale; after 20 s d Cn / d Cn0 has decayed below 1e-300 (it must have on both
    # sides), so ALL entries are compared again at nodes where they are still representable
    assert entry_err(gs, ref["sens_final"], keep=(1, 2), tiny=1e-200) < RTOL64
    assert abs(ref["sens_final"][0, 0, 0]) < 1e-200 and np.min(np.abs(ref["sens_final"][:, 1:, :])) > 1e-12
    st = np.transpose(res.sens_trace.cpu().numpy(), (3, 0, 1, 2))  # [B, ns+1, nx, nd]
    for node in (1, 5, 33, 66):   # 30 ms .. 2 s: the Cn0 column has decayed by exp(-100) at most
        assert np.min(np.max(np.abs(ref["sens_trace"][:, node, :, 0]), axis=0)) > 1e-60
        assert entry_err(st[:, node], ref["sens_trace"][:, node], keep=(1, 2)) < RTOL64, node


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_ivp_fp32_mode_holds_over_a_long_protocol(backend):
    """fp32 mode over 20 s of the fatigue protocol (6 600 RK4 steps).  What would break it in plain float arithmetic --
    A, Tau1, Km accumulating increments of ~1e-4 of their size, and decay factors 1 - h/tau_fat that a float cannot hold --
    is handled in the kernel (double accumulators for the three fatigue states, decays applied as exact small corrections,
    fes_core.cuh): states stay at float rounding level, sensitivities well inside north_star's 1e-4."""
    kind, fat = MODELS["ding2003_with_fatigue"]
    T = 20
    stim = [s + k / 33 for s in range(0, T, 2) for k in range(33)]
    ns = 33 * T
    win, _ = fes.stim_windows(stim, ns, T, 10)
    B = 8
    rng = np.random.default_rng(0)
    params = fes.default_params(kind)[None] * np.concatenate([rng.uniform(0.8, 1.2, (B, 8)), np.ones((B, 6))], axis=1)
    x0 = np.zeros((B, 5))
    x0[:, 2], x0[:, 3], x0[:, 4] = params[:, 0], params[:, 2], params[:, 1]
    f32 = torch.float32
    p_t, x_t = torch.tensor(params.T.copy(), dtype=f32, device=backend), torch.tensor(x0.T.copy(), dtype=f32, device=backend)
    res = fes.ivp(kind, fat, p_t, x_t, torch.tensor(win, device=backend), T / ns, n_steps=10, n_shooting=ns, trace="nodes",
                  dirs=fes.FES_DIR_X0, sens_trace=True, dtype=f32)
    # the oracle integrates the inputs the library was given (rounded to float), in double
    ref = foc.ivp(kind, fat, 4, 10, ns, 10, T / ns, 0.020, 1.04, p_t.double().cpu().numpy().T.copy(),
                  x_t.double().cpu().numpy().T.copy(), win[None], None, seed_x0=True, trace_B=B, sens_trace=True)
    tr = np.transpose(res.trace.cpu().numpy(), (2, 0, 1))
    st = np.transpose(res.sens_trace.cpu().numpy(), (3, 0, 1, 2))
    for node in (17, 330, 611, 660):   # on-phase, end of an off-phase, last on-phase, end
        assert entry_err(tr[:, node], ref["trace"][:, node], keep=(1,), tiny=1e-200) < 1e-5, node          # 10 x inside 1e-4
        assert entry_err(st[:, node], ref["sens_trace"][:, node], keep=(1, 2), tiny=1e-200) < RTOL32, node


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_ivp_argument_errors(backend):
    from cocofest_b200._lib import FesLibraryError

    pb = random_problem("ding2007", B=4)
    params, x0, win, ctrl = to_lib(pb, backend)
    with pytest.raises(ValueError):
        fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], ctrl=None, n_shooting=pb["ns"])
    with pytest.raises(ValueError):
        fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], ctrl=ctrl, n_shooting=pb["ns"], dirs=fes.FES_DIR_U)
    with pytest.raises(FesLibraryError):
        fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], ctrl=ctrl, n_shooting=pb["ns"], scheme=3)
    with pytest.raises(FesLibraryError):
        fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], ctrl=ctrl, n_shooting=pb["ns"], n_steps=1000)


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", ["ding2003_with_fatigue", "hmed2018"])
@pytest.mark.parametrize("B", [1, 130])
def test_ivp_edge_sizes(backend, name, B):
    """Limits of the ABI: truncation window of 64 pulses (FesDesc.n_trunc max), 256 sub-steps per interval
    (n_steps max), a single trajectory and a batch that fills two blocks and two lanes of a third."""
    pb = random_problem(name, B=B, N=64, ns=3, freq=50.0, seed=8, shared_ctrl=(B == 1))
    params, x0, win, ctrl = to_lib(pb, backend)
    S = 256
    res = fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], n_steps=S, ctrl=ctrl, n_shooting=pb["ns"],
                  dirs=fes.FES_DIR_X0, trace="nodes")
    ref = foc.ivp(pb["kind"], pb["fat"], 4, pb["N"], pb["ns"], S, pb["dt"], pb["tauc"], fo.R0_KM_RELATIONSHIP,
                  pb["params"], pb["x0"], oracle_win(pb), pb["u"], seed_x0=True)
    assert rel_err(res.x_final.cpu().numpy().T, ref["final"]) < RTOL64
    gs = np.transpose(res.sens_final.cpu().numpy(), (2, 0, 1))
    assert entry_err(gs, ref["sens_final"], keep=(1, 2)) < RTOL64
    assert res.trace.shape == (pb["ns"] + 1, pb["nx"], B)
    from cocofest_b200._lib import FesLibraryError
    with pytest.raises(FesLibraryError):  # one past the limits
        fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], n_steps=257, ctrl=ctrl, n_shooting=pb["ns"])


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_ivp_window_of_pads_only_is_pure_decay(backend):
    """Before the first pulse the window holds only the -10 000 000 pads (ding2003.py:394-397): every exp
    underflows to 0, Cn decays as exp(-t/tauc) under RK4 and F relaxes -- no NaN, same numbers as the oracle."""
    kind, fat = MODELS["ding2003_with_fatigue"]
    ns, dt, N, S = 4, 0.05, 5, 10
    win = np.full((ns + 1, N), float(fes.PAD_TIME))
    p = fo.default_params(kind)
    x0 = np.array([[0.5, 80.0, 2800.0, 0.055, 0.12]])
    res = fes.ivp(kind, fat, torch.tensor(p, device=backend), torch.tensor(x0.T.copy(), device=backend),
                  torch.tensor(win, device=backend), dt, n_steps=S, n_shooting=ns, trace="nodes")
    ref = foc.ivp(kind, fat, 4, N, ns, S, dt, fo.default_tauc(kind), fo.R0_KM_RELATIONSHIP, p[None], x0, win[None])
    got = res.x_final.cpu().numpy().T
    assert np.all(np.isfinite(got)) and rel_err(got, ref["final"]) < RTOL64
    # Cn obeys the linear test equation: one RK4 step multiplies it by the degree-4 Taylor polynomial of exp(-h/tauc)
    z = -(dt / S) / fo.default_tauc(kind)
    amp = (1 + z + z * z / 2 + z ** 3 / 6 + z ** 4 / 24) ** (ns * S)
    assert abs(got[0, 0] / (0.5 * amp) - 1) < 1e-12


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
def test_ivp_out_buffers_are_reused(backend):
    pb = random_problem("ding2003_with_fatigue", B=9, seed=2)
    params, x0, win, ctrl = to_lib(pb, backend)
    kw = dict(n_shooting=pb["ns"], dirs=fes.FES_DIR_X0, trace="nodes", sens_trace=True)
    a = fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], **kw)
    ref = [t.clone() for t in (a.x_final, a.trace, a.sens_final, a.sens_trace)]
    ptrs = [t.data_ptr() for t in (a.x_final, a.trace, a.sens_final, a.sens_trace)]
    for t in (a.x_final, a.trace, a.sens_final, a.sens_trace):
        t.zero_()
    b = fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], out=a, **kw)
    got = (b.x_final, b.trace, b.sens_final, b.sens_trace)
    assert [t.data_ptr() for t in got] == ptrs
    assert all(torch.equal(g, r) for g, r in zip(got, ref))
    with pytest.raises(ValueError, match="out: expected"):
        fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], out=a, **{**kw, "n_shooting": pb["ns"] - 1})


@pytest.mark.parametrize("backend", BACKENDS, indirect=True)
@pytest.mark.parametrize("name", ["ding2003_with_fatigue", "ding2007", "hmed2018_with_fatigue"])
def test_ivp_plan_equals_one_shot_calls(backend, name):
    """IvpPlan = ivp() with the set-up done once: same bits on every call, for new initial states / parameters / controls."""
    pb = random_problem(name, B=11, seed=6, shared_ctrl=False)
    params, x0, win, ctrl = to_lib(pb, backend)
    kw = dict(n_steps=5, n_shooting=pb["ns"], dirs=fes.FES_DIR_X0, trace="nodes", sens_trace=True, trace_batch=4)
    plan = fes.IvpPlan(pb["kind"], pb["fat"], params, x0, win, pb["dt"], ctrl=ctrl, **kw)
    for it in range(3):
        x_it, p_it = (x0 * (1.0 + 0.01 * it)).contiguous(), (params * (1.0 + 0.002 * it)).contiguous()
        u_it = None if ctrl is None else (ctrl * (1.0 - 0.01 * it)).contiguous()
        got = plan(x_it, p_it, u_it)
        ref = fes.ivp(pb["kind"], pb["fat"], p_it, x_it, win, pb["dt"], ctrl=u_it, **kw)
        for a, b in ((got.x_final, ref.x_final), (got.trace, ref.trace), (got.sens_final, ref.sens_final), (got.sens_trace, ref.sens_trace)):
            assert torch.equal(a, b)
        assert got.columns == ref.columns
    with pytest.raises(ValueError, match="x0: expected"):
        plan(x0[:, :-1].contiguous())
