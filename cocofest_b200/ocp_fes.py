"""Multiple-shooting side of the path: ``OcpFes.prepare_n_shooting`` and the batched evaluation of the
continuity defects and their Jacobian that Ipopt consumes.

Reference: cocofest/optimization/fes_ocp.py:192-222 (grid), bioptim limits/penalty.py:1056-1105
(g_k = x_{k+1} - Phi_k(x_k, u_k, p)), optimization/optimization_vector.py:121-141
(v = [dt | X_0..X_ns | U_0..U_{ns-1} | P]).  In the reference the defect and its derivatives are a closed
casadi graph differentiated by AD on every Ipopt iteration; ``FesShooting`` evaluates the same numbers for B
scenarios x ns intervals in one kernel launch per direction tile and lays the Jacobian out as COO triplets in
bioptim's variable ordering.  NLP assembly and Ipopt stay on the host, untouched.
"""
from __future__ import annotations

import numpy as np
import torch

from . import fes as _fes
from ._lib import FES_DING2007, FES_HMED2018
from .ode_solver import OdeSolver, scheme_and_steps


class OcpFes:
    """Only the static pieces of the reference class that the hot path needs."""

    @staticmethod
    def prepare_n_shooting(stim_time, final_time):
        """Number of shooting intervals such that every pulse falls on a node (fes_ocp.py:192-222).
        Warns like the reference when the grid becomes large."""
        n_shooting = _fes.prepare_n_shooting(stim_time, final_time)
        if n_shooting >= 1000:
            print(f"Warning: The number of shooting nodes is very high n = {n_shooting}.\n"
                  "The optimization might be long, consider using stimulation time with even spacing (common frequency).")
        return n_shooting


class FesShooting:
    """Batched multiple-shooting defects and Jacobians of one FES model on a fixed node grid.

    model: a cocofest_b200 model (its ``stim_time`` / ``previous_stim`` define the per-node windows)
    n_shooting, final_time: the grid (defaults: ``prepare_n_shooting``)
    ode_solver: OdeSolver.RK1/RK2/RK4 (OcpFes defaults to RK1(10), fes_ocp.py:120; examples pass RK4(10))
    """

    def __init__(self, model, final_time, n_shooting=None, ode_solver=None, device=None, dtype=torch.float64,
                 t_start=0.0):
        self.model = model
        self.final_time = final_time
        self.n_shooting = n_shooting if n_shooting is not None else OcpFes.prepare_n_shooting(model.stim_time, final_time)
        self.dt = final_time / self.n_shooting
        self.scheme, self.n_steps = scheme_and_steps(ode_solver if ode_solver is not None else OdeSolver.RK1(10))
        self.device = torch.device(device) if device is not None else torch.device("cuda")
        self.dtype = dtype
        self.t_start = t_start
        data, self.stim_idx_at_node_list = model.get_numerical_data_time_series(self.n_shooting, final_time)
        self.window = torch.as_tensor(np.ascontiguousarray(data["stim_time"][:, 0, :].T), device=self.device)
        self.N = self.window.shape[1]
        self.nx = model.nb_state
        self.nu = {FES_DING2007: 1, FES_HMED2018: self.N}.get(model.kind, 0)

    # ---- evaluation ------------------------------------------------------------------------------------
    def _params(self, params):
        return torch.as_tensor(self.model.param_vector() if params is None else params).to(self.device)

    def evaluate(self, X, U=None, params=None, want=("x", "u"), defects=True, endpoints=False, force_factor=None) -> _fes.ShootResult:
        """X [ns+1, nx, B]; U [ns, nu(, B)]; params None | [14(, B)].  Returns fes.ShootResult.
        force_factor [ns(, B)]: ``fl * fv + fp`` of every interval (FES-MSK coupling, see ``fes.shoot``); adds ``Jc``."""
        return _fes.shoot(self.model.kind, self.model.with_fatigue, self._params(params),
                          torch.as_tensor(X).to(self.device), self.window, self.dt, n_steps=self.n_steps,
                          ctrl=None if U is None else torch.as_tensor(U).to(self.device), scheme=self.scheme,
                          want=want, t_start=self.t_start, tauc=self.model.tauc, r0_km=self.model.r0_km_relationship,
                          defects=defects, endpoints=endpoints, dtype=self.dtype,
                          force_factor=None if force_factor is None else torch.as_tensor(force_factor).to(self.device))

    def plan(self, batch, U=None, params=None, want=("x", "u"), defects=True, endpoints=False, force_factor=False) -> _fes.ShootPlan:
        """The same evaluation set up once for a solver loop (``fes.ShootPlan``): ``p = shooting.plan(B, U0); res = p(X, U)``
        per iteration.  ``U`` / ``params`` are templates of the shapes the calls will use; the plan's result tensors are
        reused from call to call.  force_factor=True: ``U`` is [ns, nu + 1, B], the force factor in its last row."""
        return _fes.ShootPlan(self.model.kind, self.model.with_fatigue, self._params(params), self.window, self.dt,
                              self.n_shooting, batch, n_steps=self.n_steps,
                              ctrl=None if U is None else torch.as_tensor(U).to(self.device), scheme=self.scheme, want=want,
                              t_start=self.t_start, tauc=self.model.tauc, r0_km=self.model.r0_km_relationship,
                              defects=defects, endpoints=endpoints, dtype=self.dtype, force_factor=force_factor)

    # ---- second derivatives (for an exact-Hessian solver) ---------------------------------------------
    # typical magnitudes of (Cn, F, A, Tau1, Km) and of the controls: floors of the finite-difference steps
    _STATE_SCALE = (0.1, 10.0, 1000.0, 0.05, 0.1)

    def second_derivatives(self, X, U=None, params=None, rel_step=1e-5, method="exact", force_factor=None):
        """d2 Phi_k / d z_k d z_k for z_k = (x_k, u_k): tensor [ns, nx, nz, nz, B], nz = nx + nu (+ 1 with force_factor).

        method="exact" (default): the second-order tangent kernel (``fes.shoot_hessian`` -> fes_shoot_hess_f64), one launch,
        exact to rounding.  method="fd": the route below, kept as an independent check and for float32 plans.

        The reference's Ipopt runs on casadi's exact Hessian by default (bioptim interfaces/ipopt_options.py:88).  The
        kernels propagate first-order tangents only; this evaluates the second derivatives as CENTRAL DIFFERENCES OF THE
        EXACT JACOBIAN blocks -- 2 nz evaluations of one ``ShootPlan`` (every interval of every scenario is perturbed in the
        same call: interval k depends on x_k, u_k only) -- and symmetrises.  With the Jacobian exact to rounding the error
        is O(rel_step^2) truncation + O(1e-16 / rel_step) rounding: ~1e-10 relative at the default step (measured against
        complex-step differentiation of the oracle, tests/test_second_derivatives.py).  Not a second-order tangent
        propagation; DESIGN.md section 8 lists that as open."""
        if method not in ("exact", "fd"):
            raise ValueError("method must be 'exact' or 'fd'")
        if method == "exact":
            return self.hessian(X, U, params, force_factor=force_factor).full()
        if force_factor is not None:
            raise ValueError("the finite-difference route does not carry the force factor: use method='exact'")
        X = torch.as_tensor(X).to(self.device).contiguous()
        ns, nx, nu = self.n_shooting, self.nx, self.nu
        B = X.shape[-1]
        if nu:
            U = torch.as_tensor(U).to(self.device)
            if U.dim() == 2:
                U = U[:, :, None].expand(ns, nu, B)
            U = U.contiguous()
        plan = self.plan(B, U, params, want=("x", "u") if nu else ("x",), defects=False, endpoints=False)
        nz = nx + nu
        out = torch.empty((ns, nx, nz, nz, B), dtype=X.dtype, device=self.device)
        scale = torch.tensor(self._STATE_SCALE[:nx], dtype=X.dtype, device=self.device)

        def jac(Xp, Up):
            r = plan(Xp, Up)
            return r.Jx.clone() if not nu else torch.cat([r.Jx, r.Ju], dim=2)      # [ns, nx, nz, B]

        for d in range(nz):
            if d < nx:
                h = rel_step * torch.maximum(X[:ns, d].abs(), scale[d].expand_as(X[:ns, d]))   # [ns, B]
                Xp, Xm = X.clone(), X.clone()
                Xp[:ns, d] += h
                Xm[:ns, d] -= h
                h = (Xp[:ns, d] - Xm[:ns, d]) * 0.5        # the step that was really taken
                out[:, :, :, d] = (jac(Xp, U) - jac(Xm, U)) / (2.0 * h[:, None, None, :])
            else:
                j = d - nx
                h = rel_step * U[:, j].abs().clamp_min(1e-30)
                Up, Um = U.clone(), U.clone()
                Up[:, j] += h
                Um[:, j] -= h
                h = (Up[:, j] - Um[:, j]) * 0.5
                out[:, :, :, d] = (jac(X, Up) - jac(X, Um)) / (2.0 * h[:, None, None, :])
        return 0.5 * (out + out.transpose(2, 3))

    def hessian(self, X, U=None, params=None, multipliers=None, second=True, force_factor=None) -> _fes.HessResult:
        """Exact second derivatives in packed form (``fes.HessResult``): D2 [ns, nx, npairs, B] and, with ``multipliers``
        [ns, nx, B], H [ns, npairs, B] = sum_r lambda[k, r] d2 Phi_k[r] -- one launch per solver iteration
        (``second=False`` skips D2).  ``.full()`` / ``.full(res.H)`` give the symmetric blocks."""
        X = torch.as_tensor(X).to(self.device)
        return _fes.shoot_hessian(self.model.kind, self.model.with_fatigue, self._params(params), X, self.window[:self.n_shooting],
                                  self.dt, n_steps=self.n_steps, ctrl=U, scheme=self.scheme, multipliers=multipliers,
                                  second=second, t_start=self.t_start, tauc=self.model.tauc,
                                  r0_km=self.model.r0_km_relationship, force_factor=force_factor)

    @staticmethod
    def lagrangian_hessian(multipliers, second):
        """sum_j lambda[k, j] d2 Phi_kj / dz dz: [ns, nz, nz, B] from multipliers [ns, nx, B] and ``second_derivatives``.
        (The continuity rows are g_k = x_{k+1} - Phi_k: their contribution to the Lagrangian Hessian is minus this.)"""
        return torch.einsum("kjb,kjdeb->kdeb", multipliers, second)

    def defects(self, X, U=None, params=None) -> torch.Tensor:
        """g [ns, nx, B]: the continuity constraint values (penalty.py:1092-1099)."""
        return self.evaluate(X, U, params, want=()).G

    # ---- Jacobian in bioptim's variable ordering ------------------------------------------------------
    def variable_layout(self, n_params=0):
        """Offsets of dt, X, U, P inside bioptim's decision vector (optimization_vector.py:121-141)."""
        ns, nx, nu = self.n_shooting, self.nx, self.nu
        off_x = 1
        off_u = off_x + (ns + 1) * nx
        off_p = off_u + ns * nu
        return {"dt": 0, "x": off_x, "u": off_u, "p": off_p, "size": off_p + n_params}

    def jacobian_coo(self, res: _fes.ShootResult):
        """(rows, cols, vals[nnz, B]) of d g / d v for the continuity rows (row = k*nx + i).

        d g_k/d x_k = -dPhi/dx, d g_k/d x_{k+1} = I, d g_k/d u_k = -dPhi/du, d g_k/d p = -dPhi/dp.  The dt column
        is not produced: every cocofest FES problem fixes the phase time (fes_ocp.py phase_time=final_time)."""
        ns, nx, nu = self.n_shooting, self.nx, self.nu
        n_p = 0 if res.Jp is None else res.Jp.shape[2]
        lay = self.variable_layout(n_p)
        k = torch.arange(ns).view(ns, 1, 1)
        i = torch.arange(nx).view(1, nx, 1)
        rows, cols, vals = [], [], []

        def add(block, col_of_j):
            nj = block.shape[2]
            j = torch.arange(nj).view(1, 1, nj)
            rows.append((k * nx + i + 0 * j).reshape(-1))
            cols.append((col_of_j(k, j) + 0 * i).reshape(-1))
            vals.append(-block.reshape(ns * nx * nj, -1))

        if res.Jx is not None:
            add(res.Jx, lambda k, j: lay["x"] + k * nx + j)
        B = res.G.shape[-1] if res.G is not None else res.Jx.shape[-1]
        rows.append((k * nx + i).reshape(-1))
        cols.append((lay["x"] + (k + 1) * nx + i).reshape(-1))
        ref = res.Jx if res.Jx is not None else res.G
        vals.append(torch.ones((ns * nx, B), dtype=ref.dtype, device=ref.device))
        if res.Ju is not None:
            add(res.Ju, lambda k, j: lay["u"] + k * nu + j)
        if res.Jp is not None:
            add(res.Jp, lambda k, j: lay["p"] + 0 * k + j)
        dev = ref.device
        return torch.cat(rows).to(dev), torch.cat(cols).to(dev), torch.cat(vals, dim=0)

    def jacobian_dense(self, res: _fes.ShootResult, scenario=0) -> torch.Tensor:
        """Dense [ns*nx, len(v)] Jacobian of one scenario (debug / small problems)."""
        rows, cols, vals = self.jacobian_coo(res)
        n_p = 0 if res.Jp is None else res.Jp.shape[2]
        out = torch.zeros((self.n_shooting * self.nx, self.variable_layout(n_p)["size"]), dtype=vals.dtype, device=vals.device)
        out[rows, cols] = vals[:, scenario]
        return out

    # ---- controls helper: the Hmed sliding-window constraint (custom_constraints.py:101-119) ----------
    def pulse_intensity_controls(self, pulse_intensity, min_intensity=None):
        """Per-node control vectors consistent with per-pulse intensities: slot j of node k holds the intensity
        of the pulse at that slot of the node's window; slots before the first pulse get ``min_intensity``
        (default: the model's min_pulse_intensity, as the reference constraint does)."""
        lo = self.model.min_pulse_intensity() if min_intensity is None else min_intensity
        pi = torch.as_tensor(pulse_intensity, dtype=torch.float64)
        batched = pi.dim() == 2
        ns, N = self.n_shooting, self.N
        U = torch.full((ns, N) + ((pi.shape[1],) if batched else ()), float(lo), dtype=torch.float64)
        for k in range(ns):
            idx = self.stim_idx_at_node_list[k]
            for s, src in enumerate(idx):
                U[k, N - len(idx) + s] = pi[src]
        return U.to(self.device)


class FesNmpcBatch:
    """Window carry-over and warm-start shifting of the cyclic NMPC for B concurrent scenarios, on device tensors.

    Reference (one scenario, Python objects rebuilt every cycle): ``FesNmpc.update_stim``
    (optimization/fes_nmpc.py:69-85: the last ``sum_stim_truncation`` pulses of the horizon, shifted back by the
    phase time, become ``previous_stim`` of a NEW model whose window is rebuilt; see ``update_stim`` for the one
    deliberate deviation, the shift under a partial advance),
    ``advance_window_bounds_states`` / ``advance_window_initial_guess_states`` (:44-52, bioptim
    MultiCyclicNonlinearModelPredictiveControl: the state at frame ``n_cycles_to_advance * cycle_len`` becomes the
    fixed initial state, the remaining frames shift to the front and the tail repeats the last cycle) and
    ``initialize_frames_and_parameters`` (:33-42).  Here the horizon's window table is built once per distinct
    ``previous_stim`` (the first horizon uses the model's own, every later one the carried pulses -- identical from
    the second horizon on because the stimulation train is periodic), and the shifts are index operations on
    ``[frames, nx, B]`` / ``[frames, nu, B]`` tensors: no Python object per scenario, no rebuild per cycle.
    """

    def __init__(self, model, cycle_duration, n_cycles_simultaneous=1, n_cycles_to_advance=1, ode_solver=None,
                 device=None, dtype=torch.float64):
        if n_cycles_to_advance < 1 or n_cycles_to_advance > n_cycles_simultaneous:
            raise ValueError("n_cycles_to_advance must lie in 1..n_cycles_simultaneous")
        self.model = model
        self.cycle_duration = cycle_duration
        self.n_cycles_simultaneous = n_cycles_simultaneous
        self.n_cycles_to_advance = n_cycles_to_advance
        self.final_time = cycle_duration * n_cycles_simultaneous
        self.n_shooting = OcpFes.prepare_n_shooting(model.stim_time, self.final_time)
        if self.n_shooting % n_cycles_simultaneous:
            raise ValueError("the stimulation train must repeat identically in every simultaneous cycle")
        self.cycle_len = self.n_shooting // n_cycles_simultaneous
        self.device = torch.device(device) if device is not None else torch.device("cuda")
        solver = ode_solver if ode_solver is not None else OdeSolver.RK4(1)  # fes_ocp_nmpc_cyclic.py:40
        self.first = FesShooting(model, self.final_time, self.n_shooting, solver, device=self.device, dtype=dtype)
        self.shooting = self.first
        self.horizon = 0

    def update_stim(self, pulse_intensity=None):
        """fes_nmpc.py:69-85 -- carried pulses of the finished horizon become previous_stim of the next.

        Deviation from the reference, on purpose: the reference shifts the last ``sum_stim_truncation`` pulses back by the
        WHOLE phase time (``phase_time[0]`` = ``cycle_duration * n_cycles_simultaneous``, fes_nmpc.py:72), which is the
        start of the next horizon only when ``n_cycles_to_advance == n_cycles_simultaneous``.  Here the shift is the time
        the window really advances, ``cycle_duration * n_cycles_to_advance``, and every pulse that then lies before the
        new t = 0 is a candidate (the last ``sum_stim_truncation`` of them are kept): with a partial advance the pulses
        of the cycles that stay in the horizon are not carried twice, and closed-loop integration over the shifted
        horizons equals one long integration (tests/test_nmpc_batch.py).  The two rules agree for a full advance.

        The next model is a copy of the current one (tauc, r0_km_relationship, names and every parameter are kept), not
        a fresh ``DingModelPulseWidthFrequencyWithFatigue`` as the reference hard-codes.  Pulse-intensity models need
        the intensities of the carried pulses (hmed2018.py:43-47): pass ``pulse_intensity`` = one value per pulse of
        ``stim_time`` (the solved controls); the model's own previous intensities are carried with their pulses."""
        import copy

        N = self.model.sum_stim_truncation
        shift = self.cycle_duration * self.n_cycles_to_advance
        prev_t = list(self.model.previous_stim["time"])
        stim_t = list(self.model.stim_time)
        is_hmed = self.model.kind == _fes.FES_HMED2018
        if is_hmed:
            if pulse_intensity is None or len(pulse_intensity) != len(stim_t):
                raise ValueError("pulse-intensity models carry the intensity of every carried pulse: pass pulse_intensity "
                                 "with one value per pulse of stim_time (hmed2018.py:43-47)")
            prev_i = list(self.model.previous_stim.get("pulse_intensity", []))
            prev_i = [self.model.min_pulse_intensity()] * (len(prev_t) - len(prev_i)) + prev_i
            inten = prev_i + [float(v) for v in pulse_intensity]
        # pulses of the current horizon (and its own previous ones) that precede the next horizon's start
        pool = [(t, k) for k, t in enumerate(prev_t + stim_t) if t > _fes.PAD_TIME]
        keep = [(t - shift, k) for t, k in pool if t - shift < 0][-N:]
        carried = [t for t, _ in keep]
        nxt = copy.copy(self.model)
        nxt.stim_time = list(stim_t)
        nxt.previous_stim = {"time": list(carried)}
        if is_hmed:
            nxt.previous_stim["pulse_intensity"] = [inten[k] for _, k in keep]
        nxt.all_stim = list(carried) + list(stim_t)
        self.model = nxt
        self.shooting = FesShooting(nxt, self.final_time, self.n_shooting, OdeSolverLike(self.first.scheme, self.first.n_steps),
                                    device=self.device, dtype=self.first.dtype)
        self.horizon += 1
        return carried

    def advance(self, X, U=None):
        """Shifts the solved horizon to warm-start the next one.

        X [ns+1, nx, B], U [ns, nu, B] -> (x_start [nx, B], X_guess, U_guess): frame ``n_cycles_to_advance * cycle_len``
        becomes node 0 (its value is also returned as the fixed initial state); the tail repeats the last cycle."""
        adv = self.n_cycles_to_advance * self.cycle_len
        ns = self.n_shooting
        idx = torch.arange(ns + 1, device=X.device) + adv
        over = idx > ns
        idx = torch.where(over, idx - self.cycle_len * torch.div(idx - ns + self.cycle_len - 1, self.cycle_len, rounding_mode="floor"), idx)
        Xn = X.index_select(0, idx)
        Un = None
        if U is not None:
            iu = torch.arange(ns, device=U.device) + adv
            ou = iu > ns - 1
            iu = torch.where(ou, iu - self.cycle_len * torch.div(iu - ns + self.cycle_len, self.cycle_len, rounding_mode="floor"), iu)
            Un = U.index_select(0, iu)
        return Xn[0].clone(), Xn, Un

    def evaluate(self, X, U=None, params=None, want=("x", "u")):
        """Continuity defects and Jacobian blocks of every scenario's current horizon (``FesShooting.evaluate``)."""
        return self.shooting.evaluate(X, U, params, want=want)


class OdeSolverLike:
    """(scheme, n_integration_steps) pair accepted by ``scheme_and_steps``."""

    def __init__(self, scheme, n_integration_steps):
        self.scheme, self.n_integration_steps = scheme, n_integration_steps
