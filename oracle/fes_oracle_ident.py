"""CPU restatement of the force-identification objective and of the bounded Levenberg-Marquardt loop
--  TEST INFRASTRUCTURE ONLY (only tests/, __graft_entry__.smoke() and bench.py's cpu legs import oracle/).

What it follows in the reference (paths relative to the cocofest tree):
  * objective: Lagrange TRACK_STATE on "F", weight 1, Node.ALL, quadratic, target = force resampled at the nodes
    (optimization/fes_identification_ocp.py:248-266; identification/identification_method.py:253-278)
    ->  cost(p) = sum_{j=0..ns} w_j (F_p(t_j) - F_data(t_j))^2,  w_j = dt
  * decision parameters and their boxes (fes_identification_ocp.py:269-299;
    identification/ding2003_force_parameter_identification.py:121-150, ding2007_...:105-148, hmed2018_...:103-160)
  * fatigue models are rejected (ding2003_force_parameter_identification.py:196-199) -> two states only.

The forward map F_p(t_j) and its parameter derivatives come from the pinned C oracle (fes_oracle.c: reference
formulation, dual-number tangents).  The reference minimises this objective with Ipopt over a multiple-shooting
transcription; the product minimises it with a bounded Levenberg-Marquardt iteration on the single-shooting
form.  `levenberg_marquardt` below restates THAT iteration with numpy linear algebra so the tests can compare
iterate for iterate; "parity unpinned" applies to the optimiser (the reference holds no golden identified
parameters), the tests therefore also check the size-independent property that data generated at known
parameters are recovered.
"""
import numpy as np

from . import fes_oracle as fo
from . import fes_oracle_c as foc

P_COLS = {
    fo.DING2003: [fo.P_A_REST, fo.P_KM_REST, fo.P_TAU1_REST, fo.P_TAU2],
    fo.DING2007: [fo.P_A_REST, fo.P_KM_REST, fo.P_TAU1_REST, fo.P_TAU2, fo.P_PD0, fo.P_PDT],
    fo.HMED2018: [fo.P_A_REST, fo.P_KM_REST, fo.P_TAU1_REST, fo.P_TAU2, fo.P_AR, fo.P_BS, fo.P_IS, fo.P_CR],
}


def force_at_node_in_ocp(time, force, n_shooting, final_time):
    """identification_method.py:253-278."""
    return np.interp(np.linspace(0, final_time, n_shooting + 1), time, force)


def node_forces(kind, params, win, u, ns, S, dt, tauc, scheme=4, x0=None, r0_km=fo.R0_KM_RELATIONSHIP, n_threads=0):
    """F at the ns+1 nodes [B, ns+1] and d F / d P-columns [B, ns+1, np] of B parameter vectors [B, 14]."""
    params = np.ascontiguousarray(params, dtype=np.float64).reshape(-1, fo.NP_FULL)
    B = params.shape[0]
    N = win.shape[-1]
    x0 = np.zeros((B, 2)) if x0 is None else np.asarray(x0, dtype=np.float64).reshape(B, 2)
    out = foc.ivp(kind, False, scheme, N, ns, S, dt, tauc, r0_km, params, x0, win, u, pdir=P_COLS[kind], trace_B=B,
                  sens_trace=True, n_threads=n_threads)
    return out["trace"][:, :, 1], out["sens_trace"][:, :, 1, :]


def cost_grad_hess(kind, params, win, u, target, ns, S, dt, tauc, weight=None, **kw):
    """cost [B], grad [B, np], Gauss-Newton matrix 2 J^T W J [B, np, np]."""
    F, dF = node_forces(kind, params, win, u, ns, S, dt, tauc, **kw)
    w = np.full(ns + 1, dt) if weight is None else np.asarray(weight, dtype=np.float64)
    r = F - np.asarray(target, dtype=np.float64).reshape(-1, ns + 1)
    cost = np.einsum("j,bj,bj->b", w, r, r)
    grad = 2.0 * np.einsum("j,bj,bjd->bd", w, r, dF)
    hess = 2.0 * np.einsum("j,bjd,bje->bde", w, dF, dF)
    return cost, grad, hess


def lm_step(g, H, lam, active):
    """Marquardt-scaled step on the active columns: (H + lam diag(H)) delta = -g; None when not positive definite."""
    n = len(g)
    on = [d for d in range(n) if active[d] and H[d, d] > 0.0]
    delta = np.zeros(n)
    if not on:
        return delta
    A = H[np.ix_(on, on)].copy()
    A[np.diag_indices(len(on))] *= 1.0 + lam
    try:
        L = np.linalg.cholesky(A)
    except np.linalg.LinAlgError:
        return None
    delta[on] = np.linalg.solve(L.T, np.linalg.solve(L, -g[on]))
    return delta


def levenberg_marquardt(kind, p0, win, u, target, ns, S, dt, tauc, identify=None, lower=None, upper=None, weight=None,
                        max_iter=50, lambda0=1e-3, ftol=1e-12, xtol=1e-12, **kw):
    """One candidate (p0 [14]).  Same control flow as the product's device loop (fes_ident.cuh, ident_candidate).
    Returns dict(params, cost, grad, hess, evaluations)."""
    cols = P_COLS[kind]
    active = [True] * len(cols) if identify is None else [c in identify for c in cols]
    lo = np.full(fo.NP_FULL, -np.inf) if lower is None else np.asarray(lower, dtype=np.float64)
    hi = np.full(fo.NP_FULL, np.inf) if upper is None else np.asarray(upper, dtype=np.float64)
    ev = lambda p: [a[0] for a in cost_grad_hess(kind, p[None], win, u, target, ns, S, dt, tauc, weight=weight, **kw)]
    p = np.asarray(p0, dtype=np.float64).copy()
    if max_iter > 0:  # the iteration starts inside the box
        for d, j in enumerate(cols):
            if active[d]:
                p[j] = min(max(p[j], lo[j]), hi[j])
    cur, g, H = ev(p)
    lam, evals = lambda0, 1
    for _ in range(max_iter):
        if not (cur > 0.0) or not (lam < 1e15):
            break
        # active set: a parameter on a bound whose gradient pushes it further out is held for this step
        free = [active[d] and not (p[j] <= lo[j] and g[d] > 0.0) and not (p[j] >= hi[j] and g[d] < 0.0)
                for d, j in enumerate(cols)]
        delta = lm_step(g, H, lam, free)
        if delta is None:
            lam *= 10.0
            continue
        pt = p.copy()
        step = 0.0
        for d, j in enumerate(cols):
            if not active[d]:   # a parameter that is not identified is not projected on the box either
                continue
            v = min(max(p[j] + delta[d], lo[j]), hi[j])
            step = max(step, abs(v - p[j]) / (abs(p[j]) + 1e-30))
            pt[j] = v
        if not (step > 0.0):
            break
        ct, gt, Ht = ev(pt)
        evals += 1
        if ct < cur:
            gain = cur - ct
            p, g, H = pt, gt, Ht
            done = gain <= ftol * cur or step <= xtol
            cur = ct
            lam = max(lam / 3.0, 1e-12)
            if done:
                break
        else:
            lam *= 4.0
    return dict(params=p, cost=cur, grad=g, hess=H, evaluations=evals)
