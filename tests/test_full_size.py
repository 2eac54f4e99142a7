"""BASELINE.json configurations at their FULL sizes on the B200 (-m gpu only): spot parity of randomly chosen
trajectories against the C oracle over the whole horizon, plus size-independent properties the domain offers
(semigroup of the flow and of its transition matrices, defect-free integrated nodes, linearity of the force in
the scaling parameter, monotone Levenberg-Marquardt costs).  Tolerance 1e-9 relative (north_star)."""
import numpy as np
import pytest
import torch

from cocofest_b200 import fes, protocols
from oracle import fes_oracle as fo
from oracle import fes_oracle_c as foc
from oracle import fes_oracle_ident as foi
from tests.problems import entry_err

pytestmark = pytest.mark.gpu
RTOL = 1e-9


def _rel(a, b):
    """Arrays laid out [picked batch, ...]: every entry of the trailing axes (node, state row, direction column) is compared
    on its own scale (tests.problems.entry_err), not on the scale of the largest entry of the block."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return entry_err(a, b, keep=tuple(range(1, b.ndim)), tiny=1e-200) if b.ndim > 1 else entry_err(a, b, keep=(), tiny=1e-200)


def test_config2_full_protocol_spot_parity_and_semigroup():
    """2^20 trajectories x 300 s (9.9e10 trajectory-steps) with d x/d x0."""
    dev = torch.device("cuda", 0)
    B = 1 << 20
    proto = protocols.fatigue_protocol(final_time=300)
    ns, dt, half = proto["ns"], proto["dt"], proto["ns"] // 2
    params, x0 = protocols.fatigue_population(B, device=dev)
    win = torch.tensor(proto["win"], device=dev)
    kw = dict(n_steps=10, dirs=fes.FES_DIR_X0)
    full = fes.ivp(fes.FES_DING2003, True, params, x0, win, dt, n_shooting=ns, trace="nodes", trace_batch=1024, **kw)
    # spot parity over the whole 300 s: 16 trajectories spread over the batch vs the reference-formulation oracle
    pick = torch.linspace(0, B - 1, 16, device=dev).long()
    sub = fes.ivp(fes.FES_DING2003, True, params[:, pick].contiguous(), x0[:, pick].contiguous(), win, dt, n_shooting=ns, **kw)
    ref = foc.ivp(0, True, 4, proto["N"], ns, 10, dt, 0.020, 1.04, params[:, pick].cpu().numpy().T.copy(),
                  x0[:, pick].cpu().numpy().T.copy(), proto["win"][None], None, seed_x0=True)
    got = full.x_final[:, pick].cpu().numpy().T
    assert _rel(got, ref["final"]) < RTOL   # per state
    assert _rel(np.transpose(full.sens_final[:, :, pick].cpu().numpy(), (2, 0, 1)), ref["sens_final"]) < RTOL
    # the same trajectory gives the same bits whatever its position in the batch / the launch it sits in
    assert torch.equal(sub.x_final, full.x_final[:, pick]) and torch.equal(sub.sens_final, full.sens_final[:, :, pick])
    # semigroup: 0 -> 150 s, restart from x(150 s) with the window rows and node times of the second half
    first = fes.ivp(fes.FES_DING2003, True, params, x0, win, dt, n_shooting=half, **kw)
    second = fes.ivp(fes.FES_DING2003, True, params, first.x_final, win[half:].contiguous(), dt, n_shooting=ns - half,
                     t_start=half * dt, **kw)
    # (node times are formed as t_start + k dt in the restart, so the window tables differ in the last bits)
    assert float(((second.x_final - full.x_final).abs() / full.x_final.abs().amax(dim=1, keepdim=True)).max()) < RTOL
    chain = torch.einsum("ijb,jkb->ikb", second.sens_final, first.sens_final)
    scale = full.sens_final.abs().amax(dim=2, keepdim=True)   # every (row, column) entry on its own scale over the batch
    live = (scale > 1e-200).expand_as(chain)                  # d Cn / d Cn0 has decayed to nothing on both sides
    assert float(((chain - full.sens_final).abs() / scale.clamp_min(1e-200))[live].max()) < RTOL
    assert float(chain[~live].abs().max()) < 1e-200 and float(full.sens_final[~live].abs().max()) < 1e-200
    # the traced nodes of the first 1024 trajectories end where the final state says
    assert torch.equal(full.trace[-1], full.x_final[:, :1024])


def test_config3_full_sweep_linearity_in_a_scale_and_spot_parity():
    """4 M Ding2007 parameter sets: F and dF/d theta at 21 nodes.  Without fatigue F is linear in a_scale."""
    dev = torch.device("cuda", 0)
    B = 1 << 22
    proto, p, x0, ctrl = protocols.ding2007_sweep(B, device=dev)
    win = torch.tensor(proto["win"], device=dev)
    ns, dt = proto["ns"], proto["dt"]
    res = fes.ivp(fes.FES_DING2007, False, p, x0, win, dt, n_steps=10, ctrl=ctrl, n_shooting=ns, dirs=fes.FES_DIR_P,
                  trace="nodes", sens_trace=True)
    F, dF = res.trace[:, 1], res.sens_trace[:, 1]                    # [21, B], [21, 6, B]
    col = res.columns.index("a_scale")
    peak = F.abs().amax(dim=0).clamp_min(1e-300)
    assert float(((dF[:, col] * p[0] - F).abs() / peak).max()) < RTOL   # dF/d a_scale = F / a_scale
    pick = torch.linspace(0, B - 1, 32, device=dev).long()
    Fr, dFr = foi.node_forces(fo.DING2007, p[:, pick].cpu().numpy().T.copy(), proto["win"][None], ctrl.cpu().numpy()[None],
                              ns, 10, dt, 0.011)
    assert _rel(F[:, pick].cpu().numpy().T, Fr) < RTOL
    got = np.transpose(dF[:, :, pick].cpu().numpy(), (2, 0, 1))
    for c in range(6):
        assert _rel(got[:, :, c], dFr[:, :, c]) < RTOL, res.columns[c]
    del res, F, dF
    p2 = p.clone()
    p2[0] *= 2.0
    res2 = fes.ivp(fes.FES_DING2007, False, p2[:, : 1 << 20].contiguous(), x0[:, : 1 << 20].contiguous(), win, dt, n_steps=10,
                   ctrl=ctrl, n_shooting=ns, trace="nodes")
    res1 = fes.ivp(fes.FES_DING2007, False, p[:, : 1 << 20].contiguous(), x0[:, : 1 << 20].contiguous(), win, dt, n_steps=10,
                   ctrl=ctrl, n_shooting=ns, trace="nodes")
    pk = res1.trace[:, 1].abs().amax(dim=0).clamp_min(1e-300)
    assert float(((res2.trace[:, 1] - 2.0 * res1.trace[:, 1]).abs() / pk).max()) < RTOL


def test_config4_full_identification_costs_and_monotone_lm():
    """256 K Hmed2018 candidates: cost / gradient spot parity, LM never returns a worse point than it was given."""
    dev = torch.device("cuda", 0)
    B = 1 << 18
    proto, p, x0, ctrl = protocols.hmed2018_candidates(B, device=dev)
    win = torch.tensor(proto["win"], device=dev)
    ns, dt = proto["ns"], proto["dt"]
    truth = fo.default_params(fo.HMED2018)
    target, _ = foi.node_forces(fo.HMED2018, truth[None], proto["win"][None], ctrl.cpu().numpy()[None], ns, 10, dt, 0.020)
    tgt = torch.tensor(target[0], device=dev)
    ev = fes.identify(fes.FES_HMED2018, p, win, dt, tgt, ctrl=ctrl, max_iter=0)
    pick = torch.linspace(0, B - 1, 32, device=dev).long()
    c, g, H = foi.cost_grad_hess(fo.HMED2018, p[:, pick].cpu().numpy().T.copy(), proto["win"][None], ctrl.cpu().numpy()[None],
                                 target, ns, 10, dt, 0.020)
    assert np.max(np.abs(ev.cost[pick].cpu().numpy() - c) / c) < RTOL
    assert _rel(ev.grad[:, pick].cpu().numpy().T, g) < RTOL
    assert _rel(np.transpose(ev.hessian()[:, :, pick].cpu().numpy(), (2, 0, 1)), H) < RTOL
    lo = {k: v[0] for k, v in protocols.HMED_BOUNDS.items()}
    hi = {k: v[1] for k, v in protocols.HMED_BOUNDS.items()}
    lm = fes.identify(fes.FES_HMED2018, p, win, dt, tgt, ctrl=ctrl, lower=lo, upper=hi, max_iter=10)
    ok = torch.isfinite(ev.cost)
    assert bool((lm.cost[ok] <= ev.cost[ok]).all())
    assert float(lm.cost[ok].median()) < 1e-3 * float(ev.cost[ok].median())
    lo_v = torch.tensor([lo.get(n, -np.inf) for n in fes.PARAM_NAMES], dtype=torch.float64, device=dev)[:, None]
    hi_v = torch.tensor([hi.get(n, np.inf) for n in fes.PARAM_NAMES], dtype=torch.float64, device=dev)[:, None]
    assert bool(((lm.params >= lo_v) & (lm.params <= hi_v)).all())
    # re-evaluating the returned points reproduces the returned costs bit for bit
    again = fes.identify(fes.FES_HMED2018, lm.params, win, dt, tgt, ctrl=ctrl, max_iter=0)
    assert torch.equal(again.cost[ok], lm.cost[ok])


def test_config5_full_nmpc_defects_of_integrated_nodes_and_spot_parity():
    """64 K scenarios x 33 intervals: G, dPhi/dx, dPhi/du."""
    dev = torch.device("cuda", 0)
    B = 1 << 16
    proto, p, X, U = protocols.nmpc_scenarios(B, device=dev)
    win = torch.tensor(proto["win"], device=dev)
    ns, dt = proto["ns"], proto["dt"]
    res = fes.shoot(fes.FES_DING2007, True, p, X, win, dt, n_steps=10, ctrl=U, want=("x", "u"), endpoints=True)
    pick = torch.linspace(0, B - 1, 16, device=dev).long()
    ref = foc.shoot(fo.DING2007, True, 4, proto["N"], ns, 10, dt, 0.0, 0.011, 1.04, p.cpu().numpy()[None],
                    np.transpose(X[:, :, pick].cpu().numpy(), (2, 0, 1)).copy(), proto["win"][None],
                    np.transpose(U[:, :, pick].cpu().numpy(), (2, 0, 1)).copy(), want_x=True, want_u=True)
    assert _rel(np.transpose(res.G[:, :, pick].cpu().numpy(), (2, 0, 1)), ref["G"]) < RTOL
    assert _rel(np.transpose(res.Jx[:, :, :, pick].cpu().numpy(), (3, 0, 1, 2)), ref["Jx"]) < RTOL
    assert _rel(np.transpose(res.Ju[:, :, :, pick].cpu().numpy(), (3, 0, 1, 2)), ref["Ju"]) < RTOL
    # G = X_{k+1} - Phi_k exactly, and nodes produced by the integrator itself are defect-free
    assert torch.equal(res.G, X[1:] - res.Phi)
    run = fes.ivp(fes.FES_DING2007, True, p, X[0].contiguous(), win, dt, n_steps=10, ctrl=U, n_shooting=ns, trace="nodes")
    g = fes.shoot(fes.FES_DING2007, True, p, run.trace, win, dt, n_steps=10, ctrl=U, want=())
    assert float((g.G.abs() / run.trace[1:].abs().amax(dim=(0, 2), keepdim=True)).max()) < 1e-13
