"""CPU oracle for the FES muscle-ODE hot path  --  TEST INFRASTRUCTURE ONLY.

This file is a plain-numpy restatement of cocofest's forward integration path in the
*reference formulation* (N-term exponential sum re-evaluated at every RHS call, R_i recomputed,
two true divisions, classical RK stages exactly as bioptim unrolls them).  It is deliberately slow
and literal.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it; nothing under ``cocofest_b200/`` does.

Parity status
-------------
* forward map: PINNED.  ``tests/test_oracle_golden.py`` checks this file against every golden vector
  the reference holds for the path (``/root/reference/tests/shard1/test_ivp.py:28-44,65-89,107-124,
  150-168`` copied into ``tests/golden/ivp_golden.json``) and against RHS / window fixtures produced by
  importing the reference's own model classes in the build container
  (``tests/golden/make_reference_fixtures.py`` -> ``tests/golden/reference_rhs.json``).
* derivatives: "parity unpinned" by the reference (no reference test inspects a Jacobian; casadi is
  not installable here).  They are established by complex-step differentiation of the pinned forward
  map (``jacobian_complex_step``), which is exact to rounding for this analytic RHS.

Reference lines followed (all under /root/reference/):
  cocofest/models/ding2003.py:145-150 (get_r0/get_lambda_i), :200-252 (exp_time_fun, ri_fun, cn_sum_fun),
      :254-272 (cn_dot_fun), :274-311 (f_dot_fun), :394-429 (window builder)
  cocofest/models/ding2003_with_fatigue.py:50-59 (constants), :138-240 (system_dynamics, fatigue rows)
  cocofest/models/ding2007.py:63-79, :121-188 (a_calculation);  ding2007_with_fatigue.py:133-241
  cocofest/models/hmed2018.py:53-63, :97-98, :169-180 (lambda_i), :312-316 (pads);  hmed2018_with_fatigue.py:124-229
  cocofest/optimization/fes_ocp.py:192-222 (prepare_n_shooting)
  cocofest/integration/ivp_fes.py:232-256 (pulse modes), :312-357 (controls per node)
  external/bioptim/bioptim/dynamics/integrator.py:300-321 (RK.dxdt), :324-365 (RK1/RK2/RK4.next_x)
  external/bioptim/bioptim/interfaces/solve_ivp_interface.py:56-69 (single shooting loop)
  external/bioptim/bioptim/optimization/optimization_vector.py:366-376 (node times = dt*node)
  external/bioptim/bioptim/limits/penalty.py:1056-1105 (continuity defect)
"""

from fractions import Fraction
from math import gcd

import numpy as np

# ---- canonical per-trajectory parameter vector (same order as include/fes_b200.h) -------------
P_A_REST, P_KM_REST, P_TAU1_REST, P_TAU2 = 0, 1, 2, 3
P_ALPHA_A, P_ALPHA_TAU1, P_ALPHA_KM, P_TAU_FAT = 4, 5, 6, 7
P_PD0, P_PDT = 8, 9
P_AR, P_BS, P_IS, P_CR = 10, 11, 12, 13
NP_FULL = 14
PARAM_NAMES = [
    "a_rest", "km_rest", "tau1_rest", "tau2", "alpha_a", "alpha_tau1", "alpha_km", "tau_fat",
    "pd0", "pdt", "ar", "bs", "Is", "cr",
]

DING2003, DING2007, HMED2018 = 0, 1, 2
RK1, RK2, RK4 = 1, 2, 4
PAD_TIME = -10000000  # ding2003.py:396
PAD_INTENSITY = 50  # hmed2018.py:315


def default_params(kind: int) -> np.ndarray:
    """Default constants of the six models (ding2003.py:51-66, ding2003_with_fatigue.py:50-59,
    ding2007.py:63-79, hmed2018.py:53-63).  Note the Python literals ``10e-2`` / ``10e-6``."""
    p = np.zeros(NP_FULL)
    p[P_A_REST], p[P_TAU1_REST], p[P_TAU2], p[P_KM_REST] = 3009, 0.050957, 0.060, 0.103
    p[P_ALPHA_A], p[P_TAU_FAT] = -4.0 * 10e-2, 127
    p[P_ALPHA_TAU1], p[P_ALPHA_KM] = 2.1 * 10e-6, 1.9 * 10e-6
    p[P_PD0], p[P_PDT] = 0.000131405, 0.000194138
    p[P_AR], p[P_BS], p[P_IS], p[P_CR] = 0.586, 0.026, 63.1, 0.833
    if kind == DING2007:
        # a_scale plays the role of A_rest (ding2007_with_fatigue.py:108,211)
        p[P_A_REST], p[P_TAU1_REST], p[P_TAU2], p[P_KM_REST] = 4920, 0.060601, 0.001, 0.137
    return p


def default_tauc(kind: int) -> float:
    return 0.011 if kind == DING2007 else 0.020  # ding2007.py:71, ding2003.py:51


R0_KM_RELATIONSHIP = 1.04  # ding2003.py:52


def rest_state(kind: int, fatigue: bool, p: np.ndarray) -> np.ndarray:
    """standard_rest_values (ding2003.py:81-87, ding2003_with_fatigue.py:74-81, ding2007_with_fatigue.py:102-108)."""
    if not fatigue:
        return np.array([0.0, 0.0])
    return np.array([0.0, 0.0, p[P_A_REST], p[P_TAU1_REST], p[P_KM_REST]])


# ---- host-side grid / window logic --------------------------------------------------------------
def prepare_n_shooting(stim_time, final_time) -> int:
    """fes_ocp.py:192-222: LCM of the denominators of t_i / T as limited-denominator fractions."""
    t_final = Fraction(final_time).limit_denominator()
    n = 1
    for t in stim_time:
        d = (Fraction(t).limit_denominator() / t_final).denominator
        n = n * d // gcd(n, d)
    return n


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


def stim_windows(stim_time, n_shooting, final_time, truncation, previous_time=None, literal_compare=False):
    """ding2003.py:394-429.  Returns (win[ns+1, N], stim_idx_at_node_list).

    ``literal_compare=True`` keeps the reference's raw ``stim_time <= i*dt`` float comparison.  The
    reference's golden vectors (test_ivp.py) correspond to the grid-exact assignment (a pulse at 0.1
    belongs to node 1 although 1*(0.3/3) = 0.09999999999999999 < 0.1), which is the default here.
    """
    prev = list(previous_time) if previous_time else []
    while len(prev) < truncation:
        prev.insert(0, PAD_TIME)
    all_stim = np.array(prev + list(stim_time), dtype=float)
    dt = final_time / n_shooting
    node_idx = []
    for i in range(n_shooting + 1):
        if literal_compare:
            hit = np.where(all_stim <= i * dt)[0]
        else:
            hit = np.where(all_stim - i * dt <= 1e-9 * dt)[0]
        node_idx.append(int(hit[-1]))
    win = np.array([all_stim[: idx + 1][-truncation:] for idx in node_idx])
    node_list = list(range(n_shooting + 1))
    # ding2003.py:416-418 (the reference indexes a list of node numbers; only valid while
    # n_stim <= n_shooting + 1, which prepare_n_shooting guarantees for on-grid pulses)
    stim_idx = [list(node_list[: max(idx - truncation + 1, 0)][-truncation:]) for idx in node_idx]
    return win, stim_idx


def ivp_controls_pulse_width(pulse_width, stim_idx, n_shooting):
    """ivp_fes.py:344-351 -> u[ns, 1]."""
    if isinstance(pulse_width, list) and len(pulse_width) != 1:
        last = [stim_idx[i][-1] for i in range(len(stim_idx) - 1)]
        return np.array([[pulse_width[j]] for j in last], dtype=float)
    pw = pulse_width[0] if isinstance(pulse_width, list) else pulse_width
    return np.full((n_shooting, 1), float(pw))


def ivp_controls_pulse_intensity(pulse_intensity, stim_idx, n_shooting, truncation):
    """ivp_fes.py:326-342 (literal index formula) -> u[ns, N]."""
    if isinstance(pulse_intensity, list) and len(pulse_intensity) != 1:
        cols = [
            [
                pulse_intensity[stim_idx[i][0]]
                if i < n_shooting - j
                else pulse_intensity[stim_idx[i][(i - (n_shooting - j)) + 1]]
                for i in range(n_shooting)
            ]
            for j in range(truncation)
        ]
        return np.array(cols, dtype=float).T.copy()
    pi = pulse_intensity[0] if isinstance(pulse_intensity, list) else pulse_intensity
    return np.full((n_shooting, truncation), float(pi))


# ---- the right-hand side, reference formulation -------------------------------------------------
def cn_sum(r0, t, win, lam, tauc):
    """ding2003.py:230-252.  win[i], lam[i] may carry trailing batch axes."""
    total = 0
    n = win.shape[0]
    for i in range(n):
        if i == 0:
            ri = 1  # ding2003.py:249: first element of the *window* always gets R = 1
        else:
            ri = 1 + (r0 - 1) * np.exp(-(win[i] - win[i - 1]) / tauc)  # ri_fun :228
        e = np.exp(-(t - win[i]) / tauc)  # exp_time_fun :213
        total = total + ri * e * lam[i]
    return total


def f_dot(cn, f, a, tau1, km, tau2, force_factor=None):
    """ding2003.py:309-311; force_factor = force_length * force_velocity + passive_force (None: the defaults 1 * 1 + 0)."""
    fd = a * (cn / (km + cn)) - (f / (tau1 + tau2 * (cn / (km + cn))))
    return fd if force_factor is None else fd * force_factor


def rhs(kind, fatigue, t, x, u, p, win, tauc, r0_km=R0_KM_RELATIONSHIP, force_factor=None):
    """dx/dt for one of the six models.

    x : [nx, ...]   states Cn, F(, A, Tau1, Km)
    u : None | [1, ...] (last_pulse_width, ding2007) | [N, ...] (pulse_intensity, hmed2018)
    p : [NP_FULL, ...]
    win : [N, ...]  stimulation window of the node (oldest first)
    """
    n = win.shape[0]
    r0 = p[P_KM_REST] + r0_km  # get_r0(self.km_rest): the parameter, never the Km state (:269-271)
    if kind == HMED2018:
        lam = [p[P_AR] * (np.tanh(p[P_BS] * (u[i] - p[P_IS])) + p[P_CR]) for i in range(n)]  # hmed2018.py:179
    else:
        lam = [1 for _ in range(n)]
    s = cn_sum(r0, t, win, lam, tauc)
    cn, f = x[0], x[1]
    cn_dot = (1 / tauc) * s - (cn / tauc)  # ding2003.py:266
    if fatigue:
        a, tau1, km = x[2], x[3], x[4]
    else:
        a, tau1, km = p[P_A_REST], p[P_TAU1_REST], p[P_KM_REST]
    if kind == DING2007:
        a_eff = a * (1 - np.exp(-(u[0] - p[P_PD0]) / p[P_PDT]))  # ding2007.py:188
    else:
        a_eff = a
    fd = f_dot(cn, f, a_eff, tau1, km, p[P_TAU2], force_factor)
    if not fatigue:
        return np.stack(np.broadcast_arrays(cn_dot, fd))
    a_dot = -(a - p[P_A_REST]) / p[P_TAU_FAT] + p[P_ALPHA_A] * f
    tau1_dot = -(tau1 - p[P_TAU1_REST]) / p[P_TAU_FAT] + p[P_ALPHA_TAU1] * f
    km_dot = -(km - p[P_KM_REST]) / p[P_TAU_FAT] + p[P_ALPHA_KM] * f
    return np.stack(np.broadcast_arrays(cn_dot, fd, a_dot, tau1_dot, km_dot))


# ---- bioptim's fixed-step integrators over one shooting interval --------------------------------
def integrate_interval(kind, fatigue, scheme, n_steps, t0, dt, x0, u, p, win, tauc, r0_km=R0_KM_RELATIONSHIP, force_factor=None):
    """integrator.py:300-365.  Returns xall [nx, n_steps+1, ...] (xall[:, 0] = x0)."""
    h = dt / n_steps
    fun = lambda tt, xx: rhs(kind, fatigue, tt, xx, u, p, win, tauc, r0_km, force_factor)
    xs = [x0]
    x = x0
    for i in range(1, n_steps + 1):
        t = t0 + h * (i - 1)
        if scheme == RK1:
            x = x + h * fun(t, x)
        elif scheme == RK2:
            k1 = fun(t, x)
            x = x + h * fun(t + h / 2, x + h / 2 * k1)
        elif scheme == RK4:
            k1 = fun(t, x)
            k2 = fun(t + h / 2, x + h / 2 * k1)
            k3 = fun(t + h / 2, x + h / 2 * k2)
            k4 = fun(t + h, x + h * k3)
            x = x + h / 6 * (k1 + 2 * k2 + 2 * k3 + k4)
        else:
            raise ValueError("scheme must be RK1, RK2 or RK4")
        xs.append(x)
    return np.stack(xs, axis=1)


def ivp(kind, fatigue, scheme, n_steps, dt, x0, u, p, win, tauc, r0_km=R0_KM_RELATIONSHIP):
    """Single shooting over all nodes (solve_ivp_interface.py:56-69 with SolutionIntegrator.OCP).

    x0 [nx, B]; u None | [ns, nu(, B)]; p [NP_FULL(, B)]; win [ns(+1), N(, B)].
    Returns xall [nx, ns*S+1, B]  (IvpFes.integrate, duplicated_times=False layout).
    """
    ns = win.shape[0] - 1 if (u is None or win.shape[0] == u.shape[0] + 1) else win.shape[0]
    if u is not None:
        ns = u.shape[0]
    x = np.asarray(x0)
    out = [x[:, None]]
    for node in range(ns):
        t0 = dt * node  # optimization_vector.py:373
        # t_span[1] = t[-1] - t[0] with t = linspace(t0, t0 + dt) (solve_ivp_interface.py:58)
        dt_node = (t0 + dt) - t0
        un = None if u is None else u[node]
        xall = integrate_interval(kind, fatigue, scheme, n_steps, t0, dt_node, x, un, p, win[node], tauc, r0_km)
        out.append(xall[:, 1:])
        x = xall[:, -1]
    return np.concatenate(out, axis=1)


def shoot(kind, fatigue, scheme, n_steps, dt, X, U, p, win, tauc, r0_km=R0_KM_RELATIONSHIP, t_start=0.0, force_factor=None):
    """Multiple-shooting end points and defects (penalty.py:1092-1099).

    X [ns+1, nx, ...]; U None | [ns, nu, ...]; win [ns(+1), N, ...]
    Returns (phi [ns, nx, ...], g [ns, nx, ...]) with g_k = x_{k+1} - Phi_k(x_k, u_k, p).
    force_factor None | [ns, ...]: fl * fv + fp of the interval (ding2003.py:309-311), held constant over it.
    """
    ns = X.shape[0] - 1
    phi = []
    for k in range(ns):
        un = None if U is None else U[k]
        fc = None if force_factor is None else force_factor[k]
        xall = integrate_interval(kind, fatigue, scheme, n_steps, t_start + dt * k, dt, X[k], un, p, win[k], tauc, r0_km, fc)
        phi.append(xall[:, -1])
    phi = np.stack(phi)
    return phi, X[1:] - phi


def jacobian_complex_step(fun, z, eps=1e-30):
    """d fun / d z by the complex-step method.  fun: complex array [n] -> complex array [m]. Returns [m, n]."""
    z = np.asarray(z, dtype=complex)
    cols = []
    for j in range(z.shape[0]):
        zz = z.copy()
        zz[j] += 1j * eps
        cols.append(np.imag(fun(zz)) / eps)
    return np.stack(cols, axis=-1)
