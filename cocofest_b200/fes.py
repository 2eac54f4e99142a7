"""Batched tensor surface of the CUDA integrator (casadi-free primary API).

Layouts are the device library's: structure of arrays, **batch index last** (innermost).

    ivp(...)    single shooting of B independent trajectories over all nodes
                (what ``IvpFes.integrate`` / bioptim ``solve_ivp_interface.py:56-69`` do one trajectory at a time)
    shoot(...)  end points, continuity defects and Jacobian blocks of B x ns independent intervals
                (what bioptim ``penalty.py:1056-1105`` + casadi AD evaluate inside Ipopt)

CUDA tensors are processed in place on the current stream; CPU tensors go through the library's
host-buffer entry points (H2D -> kernels -> D2H, still computed on the GPU).  There is no CPU path.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from fractions import Fraction
from math import gcd

import numpy as np
import torch

from . import _lib
from ._lib import (FES_CTRL_SHARED, FES_DING2003, FES_DING2007, FES_DIR_C, FES_DIR_P, FES_DIR_PF, FES_DIR_U, FES_DIR_W,
                   FES_DIR_X0, FES_HMED2018, FES_NP, FES_PARAMS_SHARED, FES_TARGET_SHARED, FES_TILE, FES_WIN_SHARED,
                   PARAM_INDEX, PARAM_NAMES, FesDesc, FesIdentBuffers, FesIdentOptions, FesIvpBuffers, FesShootBuffers)

PAD_TIME = -10000000  # ding2003.py:396
PAD_INTENSITY = 50  # hmed2018.py:315
R0_KM_RELATIONSHIP = 1.04  # ding2003.py:52
DEFAULT_TAUC = {FES_DING2003: 0.020, FES_DING2007: 0.011, FES_HMED2018: 0.020}  # ding2003.py:51, ding2007.py:71

# identifiable parameters per model kind, in the column order of the FES_DIR_P group
P_COLUMNS = {
    FES_DING2003: ("a_rest", "km_rest", "tau1_rest", "tau2"),
    FES_DING2007: ("a_scale", "km_rest", "tau1_rest", "tau2", "pd0", "pdt"),
    FES_HMED2018: ("a_rest", "km_rest", "tau1_rest", "tau2", "ar", "bs", "Is", "cr"),
}
PF_COLUMNS = ("alpha_a", "alpha_tau1", "alpha_km", "tau_fat")


# ---- host-side grid / window logic (numpy; negligible cost) ---------------
def default_params(kind: int) -> np.ndarray:
    """Default constants (ding2003.py:51-66, ding2003_with_fatigue.py:50-59, ding2007.py:63-79,
    hmed2018.py:53-63) in canonical order.  The fatigue constants keep the reference's Python literals
    ``10e-2`` / ``10e-6`` (they generate the reference's golden vectors)."""
    p = np.zeros(FES_NP)
    p[0], p[1], p[2], p[3] = 3009, 0.103, 0.050957, 0.060
    p[4], p[5], p[6], p[7] = -4.0 * 10e-2, 2.1 * 10e-6, 1.9 * 10e-6, 127
    p[8], p[9] = 0.000131405, 0.000194138
    p[10], p[11], p[12], p[13] = 0.586, 0.026, 63.1, 0.833
    if kind == FES_DING2007:
        p[0], p[1], p[2], p[3] = 4920, 0.137, 0.060601, 0.001
    return p


def prepare_n_shooting(stim_time, final_time) -> int:
    """fes_ocp.py:192-222: LCM over pulses of the denominator of t_i / T as limited-denominator fractions."""
    t_final = Fraction(final_time).limit_denominator()
    n = 1
    for t in stim_time:
        d = (Fraction(t).limit_denominator() / t_final).denominator
        n = n * d // gcd(n, d)
    return n


def stim_windows(stim_time, n_shooting, final_time, truncation, previous_time=None):
    """Per-node truncated stimulation windows (ding2003.py:394-429).

    Returns ``(win[ns+1, N] float64, stim_idx_at_node_list)``.  Pulses are assigned to nodes on the
    node grid (``t_i <= n*dt`` up to 1e-9*dt): the reference's raw float compare drops a pulse at 0.1 from
    node 1 when dt = 0.3/3, but its golden vectors correspond to the grid-exact assignment.
    """
    prev = [] if previous_time is None else list(previous_time)
    while len(prev) < truncation:
        prev.insert(0, PAD_TIME)
    all_stim = np.array(prev + list(stim_time), dtype=np.float64)
    dt = final_time / n_shooting
    nodes = np.arange(n_shooting + 1) * dt
    # last index with all_stim <= node time (+ tolerance); all_stim need not be sorted w.r.t. duplicates
    hit = all_stim[None, :] - nodes[:, None] <= 1e-9 * dt
    if not hit.any(axis=1).all():
        raise ValueError("a node precedes every stimulation of the window (no previous_stim pad?)")
    node_idx = all_stim.shape[0] - 1 - np.argmax(hit[:, ::-1], axis=1)
    win = np.stack([all_stim[: i + 1][-truncation:] for i in node_idx])
    node_list = list(range(n_shooting + 1))
    stim_idx = [list(node_list[: max(int(i) - truncation + 1, 0)][-truncation:]) for i in node_idx]
    return win, stim_idx


def pulse_mode_times(stim_time, pulse_mode: str):
    """ivp_fes.py:232-256."""
    st = list(stim_time)
    if pulse_mode == "single":
        return st
    if pulse_mode == "doublet":
        st = st + [round(t + 0.005, 3) for t in st]
    elif pulse_mode == "triplet":
        st = st + [round(t + 0.005, 3) for t in st] + [round(t + 0.01, 3) for t in st]
    else:
        raise ValueError("Pulse mode not yet implemented")
    st.sort()
    return st


# ---- tensor plumbing ---------------------------------------------------------------------------------
def _stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream if device.type == "cuda" else 0


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _prep(t, dtype, device):
    if t is None:
        return None
    t = torch.as_tensor(t)
    return t.to(device=device, dtype=dtype).contiguous()


def _empty(shape, dtype, device):
    if device.type == "cpu" and torch.cuda.is_available():
        return torch.empty(shape, dtype=dtype, pin_memory=True)
    return torch.empty(shape, dtype=dtype, device=device)


def _nx(fatigue) -> int:
    return 5 if fatigue else 2


def _nu(kind, N) -> int:
    return {FES_DING2003: 0, FES_DING2007: 1, FES_HMED2018: N}[kind]


def _desc(kind, fatigue, scheme, N, ns, S, dirs, flags, B, dt, t_start, tauc, r0_km, tile_offset=0, trace_every=0,
          trace_batch=0) -> FesDesc:
    return FesDesc(_lib.FES_ABI_VERSION, int(kind), int(bool(fatigue)), int(scheme), int(N), int(ns), int(S),
                   int(dirs), int(flags), int(tile_offset), int(trace_every), 0, int(B), int(trace_batch), float(dt),
                   float(t_start), float(DEFAULT_TAUC[kind] if tauc is None else tauc), float(r0_km))


def _common(kind, fatigue, params, win, ctrl, lead, dtype, N_expected=None, extra_ctrl_rows=0):
    """Normalises params / win / ctrl, derives (device, dtype, B, N, flags).  extra_ctrl_rows: rows of ctrl after the
    model's own controls (the force factor of a shoot call)."""
    lead = torch.as_tensor(lead)
    device = lead.device
    if dtype is None:
        dtype = lead.dtype if lead.dtype in (torch.float32, torch.float64) else torch.float64
    if dtype not in (torch.float32, torch.float64):
        raise TypeError("dtype must be torch.float64 or torch.float32")
    B = lead.shape[-1]
    flags = 0
    params = _prep(params, dtype, device)
    if params.shape[0] != FES_NP:
        raise ValueError(f"params must have {FES_NP} rows ({', '.join(PARAM_NAMES)})")
    if params.dim() == 1:
        flags |= FES_PARAMS_SHARED
    elif params.shape != (FES_NP, B):
        raise ValueError(f"params must be [{FES_NP}] or [{FES_NP}, {B}]")
    win = _prep(win, torch.float64, device)
    if win.dim() == 2:
        flags |= FES_WIN_SHARED
    elif win.dim() != 3 or win.shape[2] != B:
        raise ValueError("stim window must be [rows, N] or [rows, N, B]")
    N = win.shape[1]
    nu = _nu(kind, N) + extra_ctrl_rows
    if nu:
        if ctrl is None:
            raise ValueError("this model needs controls (last_pulse_width [ns,1] / pulse_intensity [ns,N])")
        ctrl = _prep(ctrl, dtype, device)
        if ctrl.dim() == 2:
            flags |= FES_CTRL_SHARED
        elif ctrl.dim() != 3 or ctrl.shape[2] != B:
            raise ValueError("controls must be [ns, nu] or [ns, nu, B]")
        if ctrl.shape[1] != nu:
            raise ValueError(f"controls must have {nu} rows per node for this model")
    else:
        ctrl = None
    return device, dtype, B, N, flags, params, win, ctrl


@dataclass
class IvpResult:
    x_final: torch.Tensor                 # [nx, B]
    trace: torch.Tensor | None            # [npts, nx, trace_batch]
    sens_final: torch.Tensor | None       # [nx, nd, B]
    sens_trace: torch.Tensor | None       # [ns+1, nx, nd, trace_batch]
    columns: tuple                        # names of the nd sensitivity columns
    n_shooting: int
    n_steps: int


def direction_columns(kind, fatigue, dirs, N=0, tile_offset=0):
    """Names of the sensitivity columns the library produces for a direction set."""
    cols = []
    if dirs & FES_DIR_X0:
        cols += ["Cn0", "F0", "A0", "Tau10", "Km0"][: _nx(fatigue)]
    if dirs & FES_DIR_U:
        if kind == FES_DING2007:
            cols += ["last_pulse_width"]
        elif kind == FES_HMED2018:
            cols += [f"pulse_intensity_{tile_offset + i}" for i in range(FES_TILE)]
    if dirs & FES_DIR_P:
        cols += list(P_COLUMNS[kind])
        if (dirs & FES_DIR_PF) and fatigue:
            cols += list(PF_COLUMNS)
    if dirs & FES_DIR_W:
        cols += [f"stim_time_{tile_offset + i}" for i in range(FES_TILE)]
    return tuple(cols)


def ivp(kind, fatigue, params, x0, win, dt, n_steps=10, ctrl=None, n_shooting=None, scheme=4, dirs=0, t_start=0.0,
        tauc=None, r0_km=R0_KM_RELATIONSHIP, trace="none", trace_batch=None, sens_trace=False, dtype=None,
        out: "IvpResult | None" = None) -> IvpResult:
    """Integrates B trajectories over ``n_shooting`` intervals of length ``dt`` with ``n_steps`` RK steps each.

    params [14] | [14,B];  x0 [nx,B];  win [>=ns,N] | [>=ns,N,B] (float64);  ctrl None | [ns,nu] | [ns,nu,B].
    dirs: 0 | FES_DIR_X0 | FES_DIR_P | FES_DIR_P|FES_DIR_PF  -> chained sensitivities d x(t) / d (x0 | params).
    trace: "none" | "nodes" (ns+1 points) | "steps" (ns*S+1 points, the IvpFes.integrate layout), for the
    first ``trace_batch`` trajectories (default: all).
    out: an ``IvpResult`` of an earlier call with the same shapes, whose tensors are overwritten instead of
    allocating new ones (repeated calls from a solver loop; pinned host buffers are expensive to allocate).
    """
    x0 = torch.as_tensor(x0)
    device, dtype, B, N, flags, params, win, ctrl = _common(kind, fatigue, params, win, ctrl, x0, dtype)
    nx = _nx(fatigue)
    x0 = _prep(x0, dtype, device)
    if x0.shape != (nx, B):
        raise ValueError(f"x0 must be [{nx}, B]")
    ns = n_shooting if n_shooting is not None else (ctrl.shape[0] if ctrl is not None else win.shape[0] - 1)
    if win.shape[0] < ns or (ctrl is not None and ctrl.shape[0] < ns):
        raise ValueError("stim window / controls have fewer rows than n_shooting")
    nd = _lib.lib().fes_num_dirs(kind, int(bool(fatigue)), dirs)
    if nd < 0 or dirs & (FES_DIR_U | FES_DIR_W):
        raise ValueError("ivp supports dirs in {0, X0, P, P|PF}")
    modes = {"none": 0, "nodes": n_steps, "steps": 1}
    if trace not in modes:
        raise ValueError("trace must be 'none', 'nodes' or 'steps'")
    want_trace = trace != "none" or sens_trace
    trace_every = modes[trace] if trace != "none" else (n_steps if sens_trace else 0)
    tB = (B if trace_batch is None else min(int(trace_batch), B)) if want_trace else 0
    npts = ns * n_steps + 1 if trace_every == 1 else ns + 1
    desc = _desc(kind, fatigue, scheme, N, ns, n_steps, dirs, flags, B, dt, t_start, tauc, r0_km,
                 trace_every=trace_every, trace_batch=tB)
    def buf(old, shape):
        if old is None:
            return _empty(shape, dtype, device)
        if tuple(old.shape) != tuple(shape) or old.dtype != dtype or old.device != device or not old.is_contiguous():
            raise ValueError(f"out: expected a contiguous {dtype} tensor of shape {tuple(shape)} on {device}")
        return old

    out_x = buf(out and out.x_final, (nx, B))
    out_tr = buf(out and out.trace, (npts, nx, tB)) if (trace != "none" and tB) else None
    out_sf = buf(out and out.sens_final, (nx, nd, B)) if nd else None
    out_st = buf(out and out.sens_trace, (ns + 1, nx, nd, tB)) if (nd and sens_trace and tB) else None
    L = _lib.lib()
    is32 = dtype == torch.float32
    buf = FesIvpBuffers(_ptr(params), _ptr(x0), _ptr(win), _ptr(ctrl), _ptr(out_x), _ptr(out_tr), _ptr(out_sf),
                        _ptr(out_st), None, 0)
    if device.type == "cuda":
        with torch.cuda.device(device):
            wbytes = L.fes_workspace_bytes(ctypes.byref(desc), int(is32))
            ws = torch.empty(max(int(wbytes), 1), dtype=torch.uint8, device=device)
            buf.workspace, buf.workspace_bytes = ws.data_ptr(), wbytes
            fn = L.fes_ivp_f32 if is32 else L.fes_ivp_f64
            _lib.check(fn(ctypes.byref(desc), ctypes.byref(buf), ctypes.c_void_p(_stream_ptr(device))), "fes_ivp")
            ws.record_stream(torch.cuda.current_stream(device))
    else:
        fn = L.fes_ivp_host_f32 if is32 else L.fes_ivp_host_f64
        _lib.check(fn(ctypes.byref(desc), ctypes.byref(buf), None), "fes_ivp_host")
    return IvpResult(out_x, out_tr, out_sf, out_st, direction_columns(kind, fatigue, dirs, N), ns, n_steps)


class IvpPlan:
    """A batched integration set up ONCE and run many times (a solver loop over parameters, an identification on the host,
    repeated ``IvpFes.integrate()`` calls): argument checks, output tensors, workspace and -- on the first call -- the
    window tables are reused; a call only hands over the pointers of the current initial states / parameters / controls.
    For the millisecond-sized configurations the ~50 us of host work of a one-shot ``ivp()`` are 5 % of the call.

    Results are written into the plan's OWN tensors on every call.  CUDA tensors only; inputs must keep the dtype and the
    shapes given at construction; keep the calls of one plan on one stream (the tables are built on the first call's)."""

    def __init__(self, kind, fatigue, params, x0, win, dt, n_steps=10, ctrl=None, n_shooting=None, scheme=4, dirs=0,
                 t_start=0.0, tauc=None, r0_km=R0_KM_RELATIONSHIP, trace="none", trace_batch=None, sens_trace=False,
                 dtype=None):
        x0 = torch.as_tensor(x0)
        device, dtype, B, N, flags, params, win, ctrl = _common(kind, fatigue, params, win, ctrl, x0, dtype)
        if device.type != "cuda" and not _lib.is_host_emulation():
            raise ValueError("IvpPlan works on CUDA tensors (use ivp() for host tensors)")
        nx = _nx(fatigue)
        x0 = _prep(x0, dtype, device)
        if x0.shape != (nx, B):
            raise ValueError(f"x0 must be [{nx}, B]")
        ns = n_shooting if n_shooting is not None else (ctrl.shape[0] if ctrl is not None else win.shape[0] - 1)
        if win.shape[0] < ns or (ctrl is not None and ctrl.shape[0] < ns):
            raise ValueError("stim window / controls have fewer rows than n_shooting")
        L = _lib.lib()
        nd = L.fes_num_dirs(kind, int(bool(fatigue)), dirs)
        if nd < 0 or dirs & (FES_DIR_U | FES_DIR_W):
            raise ValueError("ivp supports dirs in {0, X0, P, P|PF}")
        modes = {"none": 0, "nodes": n_steps, "steps": 1}
        if trace not in modes:
            raise ValueError("trace must be 'none', 'nodes' or 'steps'")
        want_trace = trace != "none" or sens_trace
        trace_every = modes[trace] if trace != "none" else (n_steps if sens_trace else 0)
        tB = (B if trace_batch is None else min(int(trace_batch), B)) if want_trace else 0
        npts = ns * n_steps + 1 if trace_every == 1 else ns + 1
        self.device, self.dtype, self._x0, self._params, self._ctrl, self._win = device, dtype, x0, params, ctrl, win
        self._desc = _desc(kind, fatigue, scheme, N, ns, n_steps, dirs, flags, B, dt, t_start, tauc, r0_km,
                           trace_every=trace_every, trace_batch=tB)
        new = lambda shape: torch.empty(shape, dtype=dtype, device=device)
        out_x = new((nx, B))
        out_tr = new((npts, nx, tB)) if (trace != "none" and tB) else None
        out_sf = new((nx, nd, B)) if nd else None
        out_st = new((ns + 1, nx, nd, tB)) if (nd and sens_trace and tB) else None
        self.result = IvpResult(out_x, out_tr, out_sf, out_st, direction_columns(kind, fatigue, dirs, N), ns, n_steps)
        self._is32 = dtype == torch.float32
        self._fn = L.fes_ivp_f32 if self._is32 else L.fes_ivp_f64
        with self._on_device():
            wbytes = L.fes_workspace_bytes(ctypes.byref(self._desc), int(self._is32))
            self._ws = torch.empty(max(int(wbytes), 1), dtype=torch.uint8, device=device)
        self._buf = FesIvpBuffers(_ptr(params), _ptr(x0), _ptr(win), _ptr(ctrl), _ptr(out_x), _ptr(out_tr), _ptr(out_sf),
                                  _ptr(out_st), self._ws.data_ptr(), wbytes)
        self._tables = False

    def _on_device(self):
        import contextlib

        return torch.cuda.device(self.device) if self.device.type == "cuda" else contextlib.nullcontext()

    def __call__(self, x0=None, params=None, ctrl=None) -> "IvpResult":
        """x0 [nx, B], params [14(, B)], ctrl [ns, nu(, B)]: default = the tensors given at construction."""
        def check(t, like, name):
            if t is None:
                return like
            if t.dtype != like.dtype or t.device != like.device or tuple(t.shape) != tuple(like.shape) or not t.is_contiguous():
                raise ValueError(f"{name}: expected a contiguous {like.dtype} tensor of shape {tuple(like.shape)} on {like.device}")
            return t

        x0, params = check(x0, self._x0, "x0"), check(params, self._params, "params")
        buf = self._buf
        buf.x0, buf.params = x0.data_ptr(), params.data_ptr()
        if self._ctrl is not None:
            buf.ctrl = check(ctrl, self._ctrl, "ctrl").data_ptr()
        if self._tables:
            self._desc.flags |= _lib.FES_TABLES_READY
        with self._on_device():
            _lib.check(self._fn(ctypes.byref(self._desc), ctypes.byref(buf), ctypes.c_void_p(_stream_ptr(self.device))), "fes_ivp")
        self._tables = True
        return self.result

    def graph(self, x0=None, params=None, ctrl=None) -> "GraphedPlan":
        """This call as a CUDA graph over the given (or the construction-time) tensors: see ``GraphedPlan``."""
        return GraphedPlan(self, x0 if x0 is not None else self._x0, params if params is not None else self._params,
                           ctrl if ctrl is not None else self._ctrl)


class GraphedPlan:
    """An ``IvpPlan`` / ``ShootPlan`` call captured in a CUDA graph: ``replay()`` relaunches the captured kernels with one
    driver call and no Python or argument handling in between (every pass of a multi-pass plan in the one graph).  What it
    saves is HOST time per call; measured on the B200, a single trajectory of 100 RK4 steps with d x/d x0 takes 41.0 us per
    back-to-back ``IvpPlan`` call and 41.0 us per replay (tests/test_graph_plans.py): such a call is bound by the dependent
    chain of its steps, not by the launch, and large batches hide the ~10 us of host work anyway.  It
    is here for callers whose host thread is the scarce resource (a solver loop that does its own work between calls).

    The graph holds the POINTERS of the tensors given at capture: write new inputs INTO those tensors (``X.copy_(...)``) and
    call ``replay()``; results land in the plan's own tensors as always.  Built by ``plan.graph(...)``."""

    def __init__(self, plan, *inputs):
        if plan.device.type != "cuda":
            raise ValueError("CUDA graphs need a CUDA plan")
        self.plan, self.inputs = plan, inputs
        with torch.cuda.device(plan.device):
            plan(*inputs)                      # builds the window tables (not part of the graph) and warms the kernels up
            torch.cuda.current_stream(plan.device).synchronize()
            self.cuda_graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.cuda_graph):
                self.result = plan(*inputs)    # the library launches on torch's current stream: the capturing one here

    def replay(self):
        self.cuda_graph.replay()
        return self.result

    __call__ = replay


@dataclass
class ShootResult:
    G: torch.Tensor | None      # [ns, nx, B]   g_k = x_{k+1} - Phi_k
    Phi: torch.Tensor | None    # [ns, nx, B]
    Jx: torch.Tensor | None     # [ns, nx, nx, B]      dPhi_k / dx_k
    Ju: torch.Tensor | None     # [ns, nx, nu, B]      dPhi_k / du_k
    Jp: torch.Tensor | None     # [ns, nx, np, B]      dPhi_k / d identifiable params (+ fatigue params)
    Jw: torch.Tensor | None     # [ns, nx, N, B]       dPhi_k / d window pulse times
    p_columns: tuple
    Jc: torch.Tensor | None = None   # [ns, nx, B]     dPhi_k / d force factor of the interval (force_factor calls)


def _shoot_layout(kind, fatigue, nx, nu, N, want, factor=False):
    """Column counts of the Jacobian blocks (x, u, p, w), the names of the p columns and the kernel passes
    [(dirs, tile_offset, first column)] that fill them.  factor: the call carries the force factor; the blocks are x, u and
    ONE more column d / d factor right after them (every pass writes its own columns and then that column: each pass
    overwrites the previous one's copy with the same number, the last one leaves it in place)."""
    tiles = lambda n: (n + FES_TILE - 1) // FES_TILE
    if factor:
        cu = (1 if kind == FES_DING2007 else tiles(nu) * FES_TILE) if nu else 0
        passes = [(FES_DIR_X0 | FES_DIR_U | FES_DIR_C, 0, 0)]
        for t in range(1, tiles(nu) if kind == FES_HMED2018 else 1):
            passes.append((FES_DIR_U | FES_DIR_C, t * FES_TILE, nx + t * FES_TILE))
        return nx, cu, 0, 0, (), passes
    cx = nx if "x" in want else 0
    cu = (1 if kind == FES_DING2007 else tiles(nu) * FES_TILE) if "u" in want else 0
    np_cols = P_COLUMNS[kind] + (PF_COLUMNS if "pf" in want else ()) if "p" in want else ()
    cp = len(np_cols)
    cw = tiles(N) * FES_TILE if "w" in want else 0
    passes = []
    if cx and cu:
        passes.append((FES_DIR_X0 | FES_DIR_U, 0, 0))
        for t in range(1, tiles(nu) if kind == FES_HMED2018 else 1):
            passes.append((FES_DIR_U, t * FES_TILE, cx + t * FES_TILE))
    elif cx:
        passes.append((FES_DIR_X0, 0, 0))
    elif cu:
        for t in range(tiles(nu) if kind == FES_HMED2018 else 1):
            passes.append((FES_DIR_U, t * FES_TILE, t * FES_TILE))
    if cp:
        passes.append((FES_DIR_P | (FES_DIR_PF if "pf" in want else 0), 0, cx + cu))
    for t in range(tiles(N) if cw else 0):
        passes.append((FES_DIR_W, t * FES_TILE, cx + cu + cp + t * FES_TILE))
    if not passes:
        passes.append((0, 0, 0))
    return cx, cu, cp, cw, tuple(np_cols), passes


def _normalise_want(kind, fatigue, nu, want):
    want = set(want)
    if not want <= {"x", "u", "p", "pf", "w"}:
        raise ValueError("want must be a subset of {'x','u','p','pf','w'}")
    if "pf" in want:
        if not fatigue:
            raise ValueError("'pf' (fatigue parameters) needs a fatigue model")
        want.add("p")
    if nu == 0:
        want.discard("u")
    return want


def _with_factor(ctrl, force_factor, nu, ns, B, dtype, device):
    """[ns, nu + 1, B]: the model's controls followed by the force factor of the interval."""
    f = _prep(force_factor, dtype, device)
    if f.dim() == 1:
        f = f[:, None].expand(-1, B)
    if f.dim() != 2 or f.shape[0] < ns or f.shape[1] != B:
        raise ValueError("force_factor must be [ns] or [ns, B]")
    rows = [f[:ns, None, :]]
    if nu:
        if ctrl is None:
            raise ValueError("this model needs controls (last_pulse_width [ns,1] / pulse_intensity [ns,N])")
        c = _prep(ctrl, dtype, device)
        if c.dim() == 2:
            c = c[:, :, None].expand(-1, -1, B)
        if c.dim() != 3 or c.shape[0] < ns or c.shape[1] != nu or c.shape[2] != B:
            raise ValueError("controls must be [ns, nu] or [ns, nu, B]")
        rows.insert(0, c[:ns])
    return torch.cat(rows, dim=1).contiguous()


def shoot(kind, fatigue, params, X, win, dt, n_steps=10, ctrl=None, scheme=4, want=("x",), t_start=0.0, tauc=None,
          r0_km=R0_KM_RELATIONSHIP, defects=True, endpoints=False, dtype=None, force_factor=None) -> ShootResult:
    """All intervals of all scenarios at once.

    X [ns+1, nx, B];  params [14] | [14,B];  win [>=ns,N] | [>=ns,N,B];  ctrl None | [ns,nu] | [ns,nu,B].
    want: subset of {"x", "u", "p", "pf", "w"} selecting the Jacobian blocks.
    force_factor None | [ns] | [ns,B]: ``force_length * force_velocity + passive_force`` of ``f_dot_fun``
    (ding2003.py:309-311), held constant over each interval -- the seam for a caller that integrates the joint dynamics
    of an FES-MSK model itself.  Such a call returns Jx, Ju and Jc = dPhi_k / d factor_k (want must stay within {"x", "u"}).
    """
    X = torch.as_tensor(X)
    factor = force_factor is not None
    if factor:
        Xt = _prep(X, dtype if dtype is not None else (X.dtype if X.dtype in (torch.float32, torch.float64) else torch.float64), X.device)
        ctrl = _with_factor(ctrl, force_factor, _nu(kind, torch.as_tensor(win).shape[1]), Xt.shape[0] - 1, Xt.shape[-1], Xt.dtype, Xt.device)
    device, dtype, B, N, flags, params, win, ctrl = _common(kind, fatigue, params, win, ctrl, X, dtype, extra_ctrl_rows=int(factor))
    nx, nu = _nx(fatigue), _nu(kind, N)
    X = _prep(X, dtype, device)
    ns = X.shape[0] - 1
    if X.dim() != 3 or X.shape[1] != nx or ns < 1:
        raise ValueError(f"X must be [ns+1, {nx}, B]")
    if win.shape[0] < ns or (ctrl is not None and ctrl.shape[0] < ns):
        raise ValueError("stim window / controls have fewer rows than n_shooting")
    if factor and not set(want) <= {"x", "u"}:
        raise ValueError("a force_factor call returns Jx, Ju and Jc: the parameter / pulse-time blocks are not available with it")
    want = {"x", "u"} if factor else _normalise_want(kind, fatigue, nu, want)
    cx, cu, cp, cw, np_cols, passes = _shoot_layout(kind, fatigue, nx, nu, N, want, factor)
    cols = cx + cu + cp + cw + int(factor)

    L = _lib.lib()
    is32 = dtype == torch.float32
    if device.type == "cpu" and len(passes) > 1 and torch.cuda.is_available() and not _lib.is_host_emulation():
        # host tensors, several passes: stage the inputs on the device ONCE, run every pass there on a device-resident J
        # and copy each result back once (the host entry point would move the whole J both ways at every pass)
        dev = torch.device("cuda", torch.cuda.current_device())
        up = lambda t: None if t is None else t.to(dev, non_blocking=True)
        cc, ff = ctrl, None
        if factor:  # the recursive call puts the two parts together again
            cc, ff = (ctrl[:, :nu].contiguous() if nu else None), ctrl[:, nu].contiguous()
        r = shoot(kind, fatigue, up(params), up(X), up(win), dt, n_steps=n_steps, ctrl=up(cc), scheme=scheme, want=want,
                  t_start=t_start, tauc=tauc, r0_km=r0_km, defects=defects, endpoints=endpoints, dtype=dtype, force_factor=up(ff))

        def down(t):
            if t is None:
                return None
            h = _empty(t.shape, t.dtype, device)
            h.copy_(t, non_blocking=True)
            return h

        out = ShootResult(down(r.G), down(r.Phi), down(r.Jx), down(r.Ju), down(r.Jp), down(r.Jw), r.p_columns, down(r.Jc))
        torch.cuda.current_stream(dev).synchronize()
        return out

    G = _empty((ns, nx, B), dtype, device) if defects else None
    Phi = _empty((ns, nx, B), dtype, device) if endpoints else None
    J = _empty((ns, nx, cols, B), dtype, device) if cols else None

    first = True
    ws, ws_slot = None, None
    for dirs, tile_off, col0 in passes:
        desc = _desc(kind, fatigue, scheme, N, ns, n_steps, dirs, flags, B, dt, t_start, tauc, r0_km, tile_offset=tile_off)
        buf = FesShootBuffers(_ptr(params), _ptr(X), _ptr(win), _ptr(ctrl), _ptr(G) if first else None,
                              _ptr(Phi) if first else None, _ptr(J), cols, col0, None, 0)
        if device.type == "cuda":
            with torch.cuda.device(device):
                # the window tables depend on the table layout only (per-slot weights for Hmed2018 and for pulse-time
                # columns, sums otherwise): consecutive passes with the same layout reuse the workspace as it is
                slot = kind == FES_HMED2018 or bool(dirs & FES_DIR_W)
                wbytes = L.fes_workspace_bytes(ctypes.byref(desc), int(is32))
                if ws is None or slot != ws_slot:
                    ws, ws_slot = torch.empty(max(int(wbytes), 1), dtype=torch.uint8, device=device), slot
                else:
                    desc.flags |= _lib.FES_TABLES_READY
                buf.workspace, buf.workspace_bytes = ws.data_ptr(), wbytes
                fn = L.fes_shoot_f32 if is32 else L.fes_shoot_f64
                _lib.check(fn(ctypes.byref(desc), ctypes.byref(buf), ctypes.c_void_p(_stream_ptr(device))), "fes_shoot")
                ws.record_stream(torch.cuda.current_stream(device))
        else:
            fn = L.fes_shoot_host_f32 if is32 else L.fes_shoot_host_f64
            _lib.check(fn(ctypes.byref(desc), ctypes.byref(buf), None), "fes_shoot_host")
        first = False

    o = 0
    Jx = J[:, :, o:o + cx] if cx else None
    o += cx
    Ju = J[:, :, o:o + (nu if cu else 0)] if cu else None
    o += cu
    Jp = J[:, :, o:o + cp] if cp else None
    o += cp
    Jw = J[:, :, o:o + N] if cw else None
    Jc = J[:, :, cx + cu] if factor else None
    return ShootResult(G, Phi, Jx, Ju, Jp, Jw, tuple(np_cols), Jc)


@dataclass
class HessResult:
    D2: torch.Tensor | None     # [ns, nx, npairs, B]   d2 Phi_k[r] / dz_i dz_j, pairs i <= j along the upper triangle
    H: torch.Tensor | None      # [ns, npairs, B]       sum_r multipliers[k, r] D2[k, r]
    nz: int                     # z = (x, the model's controls [, force factor])

    def pair_index(self) -> torch.Tensor:
        """[nz, nz] long tensor: the packed position of entry (i, j) (symmetric)."""
        nz = self.nz
        iu = torch.triu_indices(nz, nz)
        idx = torch.zeros((nz, nz), dtype=torch.long)
        idx[iu[0], iu[1]] = torch.arange(iu.shape[1])
        idx[iu[1], iu[0]] = torch.arange(iu.shape[1])
        return idx

    def full(self, packed=None) -> torch.Tensor:
        """Packed pairs -> full symmetric blocks: D2 [ns, nx, nz, nz, B] (default) or H [ns, nz, nz, B]."""
        t = self.D2 if packed is None else packed
        idx = self.pair_index().to(t.device)
        return t.index_select(-2, idx.reshape(-1)).reshape(*t.shape[:-2], self.nz, self.nz, t.shape[-1])


def shoot_hessian(kind, fatigue, params, X, win, dt, n_steps=10, ctrl=None, scheme=4, multipliers=None, second=True,
                  t_start=0.0, tauc=None, r0_km=R0_KM_RELATIONSHIP, force_factor=None) -> HessResult:
    """Exact second derivatives of every interval map with respect to z_k = (x_k, u_k [, factor_k]): what casadi's exact
    Hessian of the continuity rows consists of (bioptim interfaces/ipopt_options.py:88, penalty.py:1056-1105).

    Arguments as ``shoot`` (fp64).  multipliers [ns, nx, B]: also returns H = sum_r multipliers[k, r] d2 Phi_k[r]; the
    continuity rows are g_k = x_{k+1} - Phi_k, so their part of the Hessian of the Lagrangian is -H.  second=False skips
    the full tensor D2 (a solver iteration needs H only).  One kernel launch, no workspace."""
    X = torch.as_tensor(X)
    factor = force_factor is not None
    if factor:
        Xt = _prep(X, torch.float64, X.device)
        ctrl = _with_factor(ctrl, force_factor, _nu(kind, torch.as_tensor(win).shape[1]), Xt.shape[0] - 1, Xt.shape[-1], torch.float64, Xt.device)
    device, dtype, B, N, flags, params, win, ctrl = _common(kind, fatigue, params, win, ctrl, X, torch.float64, extra_ctrl_rows=int(factor))
    nx, nu = _nx(fatigue), _nu(kind, N)
    X = _prep(X, dtype, device)
    ns = X.shape[0] - 1
    if X.dim() != 3 or X.shape[1] != nx or ns < 1:
        raise ValueError(f"X must be [ns+1, {nx}, B]")
    if win.shape[0] < ns or (ctrl is not None and ctrl.shape[0] < ns):
        raise ValueError("stim window / controls have fewer rows than n_shooting")
    if ctrl is not None and ctrl.shape[0] != ns:
        ctrl = ctrl[:ns].contiguous()
    if multipliers is None and not second:
        raise ValueError("nothing to compute: pass multipliers or keep second=True")
    lam = None
    if multipliers is not None:
        lam = _prep(multipliers, dtype, device)
        if tuple(lam.shape) != (ns, nx, B):
            raise ValueError(f"multipliers must be [{ns}, {nx}, {B}]")
    nz = nx + nu + int(factor)
    npairs = nz * (nz + 1) // 2
    D2 = _empty((ns, nx, npairs, B), dtype, device) if second else None
    H = _empty((ns, npairs, B), dtype, device) if lam is not None else None
    dirs = FES_DIR_X0 | FES_DIR_U | (FES_DIR_C if factor else 0)
    desc = _desc(kind, fatigue, scheme, N, ns, n_steps, dirs, flags, B, dt, t_start, tauc, r0_km)
    buf = _lib.FesHessBuffers(_ptr(params), _ptr(X), _ptr(win), _ptr(ctrl), _ptr(lam), _ptr(D2), _ptr(H))
    L = _lib.lib()
    if device.type == "cuda":
        with torch.cuda.device(device):
            _lib.check(L.fes_shoot_hess_f64(ctypes.byref(desc), ctypes.byref(buf), ctypes.c_void_p(_stream_ptr(device))), "fes_shoot_hess")
    else:
        _lib.check(L.fes_shoot_hess_host_f64(ctypes.byref(desc), ctypes.byref(buf), None), "fes_shoot_hess_host")
    return HessResult(D2, H, nz)


class ShootPlan:
    """One multiple-shooting evaluation (defects + the requested Jacobian blocks of B scenarios x ns intervals) set up
    ONCE and called once per solver iteration -- what a built OCP is to the reference (``prepare_ocp`` builds the casadi
    graph once, Ipopt evaluates it per iteration).  Everything that does not change between iterations is done at
    construction: argument checks, the pass list, the output tensors, the workspaces and -- on the first call -- the
    window tables, which later calls reuse (``FES_TABLES_READY``).  A call then only hands the library the pointers of
    the current node states / controls / parameters and launches the passes: ~10 us of host work instead of the ~50 us
    of a one-shot ``shoot()``, which for config 5 (0.46 ms of kernel time) is 10 % of the call.

    The results are written into the plan's OWN tensors on every call (views of them are returned): copy what must
    survive the next call.  Device tensors only; inputs must have the dtype and shapes given at construction; the window
    tables are built on the stream of the first call, so keep the calls of one plan on one stream (or synchronise once).
    """

    def __init__(self, kind, fatigue, params, win, dt, n_shooting, batch, n_steps=10, ctrl=None, scheme=4, want=("x",),
                 t_start=0.0, tauc=None, r0_km=R0_KM_RELATIONSHIP, defects=True, endpoints=False, dtype=torch.float64,
                 force_factor=False):
        """force_factor=True: the calls carry the force factor fl * fv + fp of every interval (see ``shoot``) as ONE MORE ROW
        of ``ctrl`` after the model's own controls -- ctrl is [ns, nu + 1, B] -- and return Jx, Ju and Jc (want is ignored)."""
        win = torch.as_tensor(win)
        device = win.device
        if device.type != "cuda" and not _lib.is_host_emulation():   # (the test suite's host build takes host pointers)
            raise ValueError("ShootPlan works on CUDA tensors (use shoot() for host tensors)")
        lead = torch.empty((1, int(batch)), dtype=dtype, device=device)
        self.factor = bool(force_factor)
        _, _, B, N, flags, params, win, ctrl = _common(kind, fatigue, params, win, ctrl, lead, dtype, extra_ctrl_rows=int(self.factor))
        self.kind, self.fatigue, self.dtype, self.device = kind, bool(fatigue), dtype, device
        self.B, self.N, self.ns, self.nx, self.nu = B, N, int(n_shooting), _nx(fatigue), _nu(kind, N)
        if win.shape[0] < self.ns:
            raise ValueError("stim window has fewer rows than n_shooting")
        if (self.nu or self.factor) and ctrl is None:
            raise ValueError("this model needs controls: pass a template `ctrl` of the shape later calls will use")
        self._params, self._win, self._ctrl = params, win, ctrl
        if self.factor and not set(want) <= {"x", "u"}:
            raise ValueError("a force_factor plan returns Jx, Ju and Jc: the parameter / pulse-time blocks are not available with it")
        want = {"x", "u"} if self.factor else _normalise_want(kind, fatigue, self.nu, want)
        cx, cu, cp, cw, self.p_columns, passes = _shoot_layout(kind, fatigue, self.nx, self.nu, N, want, self.factor)
        self._cols = (cx, cu, cp, cw)
        cols = cx + cu + cp + cw + int(self.factor)
        ns, nx = self.ns, self.nx
        self.G = torch.empty((ns, nx, B), dtype=dtype, device=device) if defects else None
        self.Phi = torch.empty((ns, nx, B), dtype=dtype, device=device) if endpoints else None
        self.J = torch.empty((ns, nx, cols, B), dtype=dtype, device=device) if cols else None
        L = _lib.lib()
        self._is32 = dtype == torch.float32
        self._fn = L.fes_shoot_f32 if self._is32 else L.fes_shoot_f64
        self._passes, self._ws = [], {}
        first = True
        with self._on_device():
            for dirs, tile_off, col0 in passes:
                desc = _desc(kind, fatigue, scheme, N, ns, n_steps, dirs, flags, B, dt, t_start, tauc, r0_km, tile_offset=tile_off)
                slot = kind == FES_HMED2018 or bool(dirs & FES_DIR_W)   # table layout: per-slot weights or sums only
                wbytes = L.fes_workspace_bytes(ctypes.byref(desc), int(self._is32))
                if slot not in self._ws:
                    self._ws[slot] = [torch.empty(max(int(wbytes), 1), dtype=torch.uint8, device=device), False]
                buf = FesShootBuffers(_ptr(params), None, _ptr(win), _ptr(ctrl), _ptr(self.G) if first else None,
                                      _ptr(self.Phi) if first else None, _ptr(self.J), cols, col0,
                                      self._ws[slot][0].data_ptr(), wbytes)
                self._passes.append((desc, buf, slot))
                first = False

    def _on_device(self):
        import contextlib

        return torch.cuda.device(self.device) if self.device.type == "cuda" else contextlib.nullcontext()

    def __call__(self, X, ctrl=None, params=None) -> "ShootResult":
        """X [ns+1, nx, B];  ctrl [ns, nu(, B)] when the model has controls (default: the tensor given at construction);
        params [14(, B)] (default: the tensor given at construction)."""
        ns, nx, B = self.ns, self.nx, self.B

        def check(t, like, name):
            if t is None:
                return like
            if t.dtype != like.dtype or t.device != like.device or tuple(t.shape) != tuple(like.shape) or not t.is_contiguous():
                raise ValueError(f"{name}: expected a contiguous {like.dtype} tensor of shape {tuple(like.shape)} on {like.device}")
            return t

        if X.dtype != self.dtype or X.device != self.device or tuple(X.shape) != (ns + 1, nx, B) or not X.is_contiguous():
            raise ValueError(f"X: expected a contiguous {self.dtype} tensor of shape {(ns + 1, nx, B)} on {self.device}")
        ctrl = check(ctrl, self._ctrl, "ctrl") if (self.nu or self.factor) else None
        params = check(params, self._params, "params")
        stream = ctypes.c_void_p(_stream_ptr(self.device))
        with self._on_device():
            for desc, buf, slot in self._passes:
                buf.X, buf.params = X.data_ptr(), params.data_ptr()
                buf.ctrl = None if ctrl is None else ctrl.data_ptr()
                state = self._ws[slot]
                if state[1]:
                    desc.flags |= _lib.FES_TABLES_READY
                _lib.check(self._fn(ctypes.byref(desc), ctypes.byref(buf), stream), "fes_shoot")
                state[1] = True
        cx, cu, cp, cw = self._cols
        J, o = self.J, 0
        Jx = J[:, :, o:o + cx] if cx else None
        o += cx
        Ju = J[:, :, o:o + self.nu] if cu else None
        o += cu
        Jp = J[:, :, o:o + cp] if cp else None
        o += cp
        Jw = J[:, :, o:o + self.N] if cw else None
        Jc = J[:, :, cx + cu] if self.factor else None
        return ShootResult(self.G, self.Phi, Jx, Ju, Jp, Jw, self.p_columns, Jc)

    def graph(self, X, ctrl=None, params=None) -> "GraphedPlan":
        """This call as a CUDA graph over ``X`` and the given (or the construction-time) controls / parameters: see
        ``GraphedPlan``.  All passes of the plan are part of the one graph."""
        return GraphedPlan(self, X, ctrl if ctrl is not None else self._ctrl, params if params is not None else self._params)


@dataclass
class IdentResult:
    params: torch.Tensor        # [14, B]  identified canonical parameter vectors
    cost: torch.Tensor          # [B]      sum_j w_j (F(t_j) - target_j)^2 at params
    grad: torch.Tensor          # [np, B]  d cost / d P-column
    hess: torch.Tensor          # [np(np+1)/2, B]  Gauss-Newton matrix 2 J^T W J, upper triangle row-major
    evaluations: torch.Tensor   # [B] int32
    columns: tuple              # names of the np P-columns

    def hessian(self) -> torch.Tensor:
        """Dense symmetric [np, np, B] view of ``hess``."""
        n = len(self.columns)
        iu = torch.triu_indices(n, n, device=self.hess.device)
        H = torch.zeros((n, n, self.hess.shape[-1]), dtype=self.hess.dtype, device=self.hess.device)
        H[iu[0], iu[1]] = self.hess
        H[iu[1], iu[0]] = self.hess
        return H

    def best(self) -> int:
        """Index of the candidate with the lowest finite cost (multi-start selection)."""
        c = torch.where(torch.isfinite(self.cost), self.cost, torch.full_like(self.cost, float("inf")))
        return int(torch.argmin(c).item())


def identify(kind, params, win, dt, target, n_steps=10, ctrl=None, identify=None, lower=None, upper=None, weight=None,
             x0=None, max_iter=50, lambda0=1e-3, ftol=1e-12, xtol=1e-12, scheme=4, t_start=0.0, tauc=None,
             r0_km=R0_KM_RELATIONSHIP, dtype=None) -> IdentResult:
    """Force-tracking cost / gradient / Gauss-Newton matrix and bounded Levenberg-Marquardt identification of
    B candidates at once (one GPU thread per candidate, the whole loop on device).

    Replaces, for the two-state models, ``OcpFesId`` + one Ipopt solve per candidate
    (optimization/fes_identification_ocp.py:248-299): cost = sum over ALL ns+1 nodes of w_j (F(t_j) - target_j)^2,
    w_j = dt by default (Lagrange TRACK_STATE on "F", Node.ALL, quadratic).

    params [14, B] (or [14], needs a per-candidate ``target``): starting points;  target [ns+1] | [ns+1, B];
    identify: names of the parameters to identify (default: all P-columns of the model, ``P_COLUMNS[kind]``);
    lower / upper: dict name -> bound, or [14] tensors;  max_iter = 0 only evaluates.
    """
    target = torch.as_tensor(target)
    params = torch.as_tensor(params)
    lead = params if params.dim() == 2 else (target if target.dim() == 2 else None)
    if lead is None:
        raise ValueError("either params [14,B] or target [ns+1,B] must carry the batch axis")
    device, dtype, B, N, flags, params, win, ctrl = _common(kind, False, params, win, ctrl, lead, dtype)
    target = _prep(target, dtype, device)
    ns = target.shape[0] - 1
    if target.dim() == 1:
        flags |= FES_TARGET_SHARED
    elif target.shape != (ns + 1, B):
        raise ValueError("target must be [ns+1] or [ns+1, B]")
    if ns < 1 or win.shape[0] < ns or (ctrl is not None and ctrl.shape[0] < ns):
        raise ValueError("stim window / controls have fewer rows than the target has intervals")
    cols = P_COLUMNS[kind]
    names = cols if identify is None else tuple(identify)
    alias = {"a_scale": "a_rest", "a_rest": "a_scale"}
    mask = 0
    for n in names:
        if n in cols:
            mask |= 1 << cols.index(n)
        elif alias.get(n) in cols:
            mask |= 1 << cols.index(alias[n])
        else:
            raise ValueError(f"{n} is not identifiable for this model; available: {cols}")

    def bounds(b, fill):
        if b is None:
            return None
        if isinstance(b, dict):
            v = torch.full((FES_NP,), fill, dtype=torch.float64)
            for k, val in b.items():
                v[PARAM_INDEX[k]] = val
            b = v
        return _prep(b, dtype, device)

    lower, upper = bounds(lower, -float("inf")), bounds(upper, float("inf"))
    weight = _prep(weight, dtype, device)
    if weight is not None and weight.shape != (ns + 1,):
        raise ValueError("weight must be [ns+1]")
    x0 = _prep(x0, dtype, device)
    if x0 is not None and x0.shape != (2, B):
        raise ValueError("x0 must be [2, B]")
    n_p = len(cols)
    desc = _desc(kind, False, scheme, N, ns, n_steps, FES_DIR_P, flags, B, dt, t_start, tauc, r0_km)
    out_p = _empty((FES_NP, B), dtype, device)
    out_c = _empty((B,), dtype, device)
    out_g = _empty((n_p, B), dtype, device)
    out_h = _empty((n_p * (n_p + 1) // 2, B), dtype, device)
    out_i = _empty((B,), torch.int32, device)
    opt = FesIdentOptions(int(max_iter), int(mask), float(lambda0), float(ftol), float(xtol))
    buf = FesIdentBuffers(_ptr(params), _ptr(x0), _ptr(win), _ptr(ctrl), _ptr(target), _ptr(weight), _ptr(lower),
                          _ptr(upper), _ptr(out_p), _ptr(out_c), _ptr(out_g), _ptr(out_h), _ptr(out_i), None, 0)
    L = _lib.lib()
    is32 = dtype == torch.float32
    if device.type == "cuda":
        with torch.cuda.device(device):
            wbytes = L.fes_workspace_bytes(ctypes.byref(desc), int(is32))
            ws = torch.empty(max(int(wbytes), 1), dtype=torch.uint8, device=device)
            buf.workspace, buf.workspace_bytes = ws.data_ptr(), wbytes
            fn = L.fes_identify_f32 if is32 else L.fes_identify_f64
            _lib.check(fn(ctypes.byref(desc), ctypes.byref(buf), ctypes.byref(opt), ctypes.c_void_p(_stream_ptr(device))),
                       "fes_identify")
            ws.record_stream(torch.cuda.current_stream(device))
    else:
        fn = L.fes_identify_host_f32 if is32 else L.fes_identify_host_f64
        _lib.check(fn(ctypes.byref(desc), ctypes.byref(buf), ctypes.byref(opt), None), "fes_identify_host")
    return IdentResult(out_p, out_c, out_g, out_h, out_i, cols)


def last_host_timing() -> dict:
    """Phases (ms) of this thread's last call that went through the library's host-buffer entry points (CPU tensors):
    staging allocation, host-to-device copies, kernels, device-to-host copies, whole call."""
    ms = (ctypes.c_double * 5)()
    _lib.check(_lib.lib().fes_last_host_timing(ms), "fes_last_host_timing")
    return dict(zip(("alloc_ms", "h2d_ms", "kernels_ms", "d2h_ms", "total_ms"), (float(v) for v in ms)))


def probe_fma_peak(is_f32=False, launches=5, device=None) -> float:
    """Measured FMA throughput of the current GPU in TFLOP/s (2 flop per FMA): the roofline denominator."""
    out = ctypes.c_double(0.0)
    dev = torch.device("cuda" if device is None else device)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().fes_probe_fma_peak(int(is_f32), int(launches), ctypes.c_void_p(_stream_ptr(dev)),
                                                 ctypes.byref(out)), "fes_probe_fma_peak")
    return out.value
