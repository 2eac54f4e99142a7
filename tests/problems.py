"""Shared problem builders for the parity tests: the same seeded inputs go to the oracle (batch-outermost
numpy layout) and to the library (SoA torch tensors, batch last)."""
import numpy as np
import torch

from oracle import fes_oracle as fo

MODELS = {
    "ding2003": (fo.DING2003, False),
    "ding2003_with_fatigue": (fo.DING2003, True),
    "ding2007": (fo.DING2007, False),
    "ding2007_with_fatigue": (fo.DING2007, True),
    "hmed2018": (fo.HMED2018, False),
    "hmed2018_with_fatigue": (fo.HMED2018, True),
}
P_COLS = {  # canonical indices of the FES_DIR_P columns, per kind
    fo.DING2003: [fo.P_A_REST, fo.P_KM_REST, fo.P_TAU1_REST, fo.P_TAU2],
    fo.DING2007: [fo.P_A_REST, fo.P_KM_REST, fo.P_TAU1_REST, fo.P_TAU2, fo.P_PD0, fo.P_PDT],
    fo.HMED2018: [fo.P_A_REST, fo.P_KM_REST, fo.P_TAU1_REST, fo.P_TAU2, fo.P_AR, fo.P_BS, fo.P_IS, fo.P_CR],
}
PF_COLS = [fo.P_ALPHA_A, fo.P_ALPHA_TAU1, fo.P_ALPHA_KM, fo.P_TAU_FAT]


def nu_of(kind, N):
    return {fo.DING2003: 0, fo.DING2007: 1, fo.HMED2018: N}[kind]


def random_problem(name, B, N=6, ns=7, freq=20.0, seed=0, shared_params=False, shared_ctrl=True, per_traj_win=False):
    """A pulse train at `freq` Hz with a few previous pulses, per-trajectory parameters +-20 %."""
    kind, fat = MODELS[name]
    rng = np.random.default_rng(seed)
    dt = 1.0 / freq
    stim = [k * dt for k in range(ns)]
    prev = [-(3 - k) * dt for k in range(3)]
    win, _ = fo.stim_windows(stim, ns, ns * dt, N, previous_time=prev)
    if per_traj_win:
        # jitter the real pulses a little, per trajectory (never past their node)
        jit = rng.uniform(0, 0.3 * dt, size=(B,) + win.shape)
        winb = np.where(win[None] > -1e6, win[None] - jit, win[None])
        winb = np.sort(winb, axis=2)
    else:
        winb = None
    p0 = fo.default_params(kind)
    params = p0[None, :] * rng.uniform(0.8, 1.2, size=(1 if shared_params else B, fo.NP_FULL))
    nx = 5 if fat else 2
    x0 = np.zeros((B, nx))
    x0[:, 0] = rng.uniform(0.0, 0.8, B)
    x0[:, 1] = rng.uniform(0.0, 150.0, B)
    if fat:
        pp = np.broadcast_to(params, (B, fo.NP_FULL))
        x0[:, 2] = pp[:, fo.P_A_REST] * rng.uniform(0.7, 1.0, B)
        x0[:, 3] = pp[:, fo.P_TAU1_REST] * rng.uniform(1.0, 1.3, B)
        x0[:, 4] = pp[:, fo.P_KM_REST] * rng.uniform(1.0, 1.3, B)
    nu = nu_of(kind, N)
    u = None
    if nu:
        shape = (1 if shared_ctrl else B, ns, nu)
        u = rng.uniform(2e-4, 6e-4, shape) if kind == fo.DING2007 else rng.uniform(30.0, 120.0, shape)
    return dict(name=name, kind=kind, fat=fat, B=B, N=N, ns=ns, dt=dt, nx=nx, nu=nu, win=win, winb=winb,
                params=params, x0=x0, u=u, tauc=fo.default_tauc(kind), shared_params=shared_params)


def to_lib(pb, device, dtype=torch.float64):
    """SoA tensors for cocofest_b200.fes (batch last)."""
    params = torch.tensor(pb["params"][0] if pb["shared_params"] else pb["params"].T.copy(), dtype=dtype, device=device)
    x0 = torch.tensor(pb["x0"].T.copy(), dtype=dtype, device=device)
    if pb["winb"] is not None:
        win = torch.tensor(np.ascontiguousarray(np.transpose(pb["winb"], (1, 2, 0))), dtype=torch.float64, device=device)
    else:
        win = torch.tensor(pb["win"], dtype=torch.float64, device=device)
    ctrl = None
    if pb["u"] is not None:
        u = pb["u"]
        ctrl = torch.tensor(u[0] if u.shape[0] == 1 else np.ascontiguousarray(np.transpose(u, (1, 2, 0))), dtype=dtype, device=device)
    return params, x0, win, ctrl


def oracle_win(pb):
    return pb["winb"] if pb["winb"] is not None else pb["win"][None]


def rel_err(a, b, floor=1e-12):
    """max |a-b| / max(|b|, floor * scale) with scale = max|b| (relative to the reference magnitude)."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    scale = max(np.max(np.abs(b)), 1e-300)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor * scale + 1e-300)))


def entry_err(got, ref, keep, tiny=1e-300):
    """Strict relative error for blocks whose entries differ by orders of magnitude (sensitivity matrices): every
    entry (the index along the axes `keep`, e.g. (state row, direction column)) is compared on ITS OWN scale, the
    largest magnitude that entry takes over the remaining axes (batch, nodes).  An entry that is structurally zero in the
    reference, or has decayed below `tiny` there, must be below `tiny` in `got` too.  Returns the worst entry's error."""
    got, ref = np.asarray(got, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    keep = tuple(k % ref.ndim for k in keep)
    red = tuple(a for a in range(ref.ndim) if a not in keep)
    scale = np.max(np.abs(ref), axis=red)
    diff = np.max(np.abs(got - ref), axis=red)
    err = np.where(scale > tiny, diff / np.maximum(scale, tiny), np.where(diff <= tiny, 0.0, np.inf))
    return float(np.max(err))
