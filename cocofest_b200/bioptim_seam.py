"""The object bioptim keeps as ``nlp.dynamics[k]``, evaluated by the CUDA library.

Reference: bioptim builds ONE ``Integrator`` per phase (``dynamics/ode_solver.py:232-249``,
``SHARED_DURING_THE_PHASE``) whose ``.function`` is the casadi Function
``integrator(t_span[2], x0[nx], u[nu], p[np], a[0], d[N]) -> (xf[nx], xall[nx, S+1])``
(``dynamics/integrator.py:82-104``); ``state_continuity`` (``limits/penalty.py:1092-1099``) and
``solve_ivp_interface`` (``interfaces/solve_ivp_interface.py:67-69``) call it, casadi differentiates it.

``FesIntegrator`` has the same call contract and needs neither casadi nor bioptim: inputs and outputs are
arrays, one column per scenario (the reference evaluates one column at a time).  Its Jacobian blocks are the
ones casadi's AD of the unrolled graph yields.  ``as_casadi_function`` / ``GpuRK4`` wrap it for a machine
that has casadi >= 3.6 and bioptim 3.2.1.  Neither is installable in this image: they raise ``ImportError`` here
(nothing is evaluated on the CPU in their place), and ``tests/test_casadi_seam.py`` executes them against strict
stand-ins of the casadi ``Callback`` contract and of ``OdeSolver.RK4`` (``tests/standins``): declarations, call,
``jac_<out>_<in>`` ordering, sparsities, block values, ``.map``.  A run against the real casadi
(``Function.jacobian()`` of the stock unrolled integrator vs these blocks) stays an open item for a machine that has it.
"""
from __future__ import annotations

import numpy as np
import torch

from . import fes as _fes
from ._lib import FES_DING2007, FES_HMED2018, PARAM_INDEX
from .ode_solver import scheme_and_steps


def _cols(a, n, name, dtype=torch.float64):
    """array-like [n] | [n, B] -> float tensor [n, B] (B = 1 for a vector); None for n = 0"""
    if n == 0:
        return None
    if a is None:
        raise ValueError(f"{name} must have {n} rows")
    t = torch.as_tensor(np.asarray(a, dtype=np.float64) if not isinstance(a, torch.Tensor) else a).to(dtype)
    if t.dim() == 1:
        t = t.reshape(-1, 1)
    if t.dim() != 2 or t.shape[0] != n:
        raise ValueError(f"{name} must be [{n}] or [{n}, B], got {tuple(t.shape)}")
    return t


class FesIntegrator:
    """``(t_span, x0, u, p, a, d) -> {"xf": [nx, B], "xall": [nx, S+1, B]}`` for one shooting interval.

    model: a cocofest_b200 model (kind, fatigue, default parameter values, tauc)
    ode_solver: OdeSolver.RK1/RK2/RK4 (or any object with ``n_integration_steps``)
    parameter_names: names of the NLP parameters ``p``, in the order the OCP declares them
        (``OcpFesId._set_parameters``, fes_identification_ocp.py:269-299); empty for ``OcpFes``.
    """

    input_names = ("t_span", "x0", "u", "p", "a", "d")      # integrator.py:134
    output_names = ("xf", "xall")

    def __init__(self, model, ode_solver, parameter_names=(), device=None, dtype=torch.float64):
        self.model = model
        self.scheme, self.n_steps = scheme_and_steps(ode_solver)
        self.kind, self.fatigue = model.kind, bool(model.with_fatigue)
        self.N = int(model.sum_stim_truncation)
        self.nx = model.nb_state
        self.nu = {FES_DING2007: 1, FES_HMED2018: self.N}.get(self.kind, 0)
        self.parameter_names = tuple(parameter_names)
        known = _fes.P_COLUMNS[self.kind] + (_fes.PF_COLUMNS if self.fatigue else ())
        for n in self.parameter_names:
            if n not in known:
                raise ValueError(f"parameter {n!r} is not identifiable for {type(model).__name__}; known: {known}")
        self._p_slot = [PARAM_INDEX[n] for n in self.parameter_names]
        self.device = torch.device(device) if device is not None else torch.device("cuda")
        self.dtype = dtype

    # ---- the Integrator surface bioptim reads -------------------------------------------------------
    @property
    def shape_xf(self):
        return (self.nx, 1)

    @property
    def shape_xall(self):
        return (self.nx, self.n_steps + 1)

    def step_times_from_dt(self, t_span):
        """Times of the columns of xall (integrator.py:247-268): t0 + dt * i / S."""
        t0, dt = float(t_span[0]), float(t_span[1])
        return t0 + dt * np.arange(self.n_steps + 1) / self.n_steps

    # ---- evaluation ----------------------------------------------------------------------------------
    def _inputs(self, x0, u, p, d):
        x0 = _cols(x0, self.nx, "x0")
        B = x0.shape[1]
        u = _cols(u, self.nu, "u")
        p = _cols(p, len(self.parameter_names), "p")
        d = _cols(d, self.N, "d")
        for name, t in (("u", u), ("p", p), ("d", d)):
            if t is not None and t.shape[1] not in (1, B):
                raise ValueError(f"{name} has {t.shape[1]} columns, x0 has {B}")
        params = torch.as_tensor(self.model.param_vector(), dtype=torch.float64)  # [14]: shared by all scenarios
        if p is not None:
            params = params.reshape(-1, 1).repeat(1, B)
            for j, slot in enumerate(self._p_slot):
                params[slot] = p[j]  # [1] broadcasts, [B] is per scenario
        win = d.T.contiguous().reshape(1, self.N) if d.shape[1] == 1 else d.reshape(1, self.N, -1)
        if win.dim() == 3 and win.shape[2] != B:
            raise ValueError("d must have one column or one per scenario")
        ctrl = None
        if u is not None:
            ctrl = u.T.contiguous().reshape(1, self.nu) if u.shape[1] == 1 else u.reshape(1, self.nu, -1)
        dev = self.device
        return (x0.to(dev), params.to(dev), win.to(dev), None if ctrl is None else ctrl.to(dev), B)

    def __call__(self, t_span, x0, u=None, p=None, a=None, d=None):
        """One interval starting at ``t_span[0]`` of length ``t_span[1]``; ``a`` (algebraic states) is unused by
        every FES model and only kept for the signature."""
        x0, params, win, ctrl, _ = self._inputs(x0, u, p, d)
        res = _fes.ivp(self.kind, self.fatigue, params, x0, win, float(t_span[1]), n_steps=self.n_steps, ctrl=ctrl,
                       n_shooting=1, scheme=self.scheme, t_start=float(t_span[0]), tauc=self.model.tauc,
                       r0_km=self.model.r0_km_relationship, trace="steps", dtype=self.dtype)
        return {"xf": res.x_final, "xall": res.trace.permute(1, 0, 2)}

    def jacobian(self, t_span, x0, u=None, p=None, a=None, d=None, stim_time=False):
        """Dense blocks of d xf / d (x0, u, p[, d]):  {"x0": [nx, nx, B], "u": [nx, nu, B], "p": [nx, np, B]}
        (+ "d": [nx, N, B], the pulse-time columns, when ``stim_time``).  Absent inputs give ``None``."""
        x0, params, win, ctrl, B = self._inputs(x0, u, p, d)
        X = torch.stack([x0, x0])  # node 1 is not read when defects are off
        want = ["x"] + (["u"] if self.nu else []) + (["w"] if stim_time else [])
        if self.parameter_names:
            want.append("pf" if any(n in _fes.PF_COLUMNS for n in self.parameter_names) else "p")
        res = _fes.shoot(self.kind, self.fatigue, params, X, win, float(t_span[1]), n_steps=self.n_steps, ctrl=ctrl,
                         scheme=self.scheme, want=want, t_start=float(t_span[0]), tauc=self.model.tauc,
                         r0_km=self.model.r0_km_relationship, defects=False, endpoints=True, dtype=self.dtype)
        out = {"xf": res.Phi[0], "x0": res.Jx[0], "u": None if res.Ju is None else res.Ju[0], "p": None}
        if self.parameter_names:
            names = ["a_rest" if n == "a_scale" else n for n in self.parameter_names]
            have = ["a_rest" if n == "a_scale" else n for n in res.p_columns]
            out["p"] = res.Jp[0][:, [have.index(n) for n in names]]
        if stim_time:
            out["d"] = res.Jw[0]
        return out

    def dense_jacobian(self, t_span, x0, u=None, p=None, a=None, d=None):
        """[nx, 2 + nx + nu + np + 0 + N] for ONE scenario, columns in the order of ``input_names``: what a casadi
        Jacobian of ``xf`` with respect to all six inputs holds.  The ``t_span`` and ``d`` columns are left zero:
        cocofest fixes the phase time and passes the window as numerical data, never as a decision variable."""
        J = self.jacobian(t_span, x0, u, p, a, d)
        npar = len(self.parameter_names)
        out = np.zeros((self.nx, 2 + self.nx + self.nu + npar + self.N))
        out[:, 2:2 + self.nx] = J["x0"][..., 0].cpu().numpy()
        if self.nu:
            out[:, 2 + self.nx:2 + self.nx + self.nu] = J["u"][..., 0].cpu().numpy()
        if npar:
            o = 2 + self.nx + self.nu
            out[:, o:o + npar] = J["p"][..., 0].cpu().numpy()
        return out


# ---- casadi / bioptim wrappers (need casadi >= 3.6 and bioptim 3.2.1; executed against tests/standins here) -----
def as_casadi_function(integrator: FesIntegrator, name="integrator"):
    """casadi Function with bioptim's integrator signature whose evaluation and first derivatives come from
    ``integrator``.  Only first-order sensitivities exist: run Ipopt with
    ``Solver.IPOPT(_hessian_approximation="limited-memory")``.

    Derivative blocks supplied: d xf / d (x0, u, p).  The blocks of ``xall`` and the columns of ``t_span`` and ``d``
    are declared structurally zero: bioptim's continuity constraint reads ``xf`` only (penalty.py:1092-1099), cocofest
    fixes the phase time and passes the stimulation window as numerical data.  A problem that puts ``xall`` or a free
    phase time into a constraint or objective must not use this wrapper."""
    try:
        import casadi
    except ImportError as e:
        raise ImportError("as_casadi_function needs casadi >= 3.6 (not installed); "
                          "use FesIntegrator / FesShooting directly") from e
    it = integrator
    sizes_in = [2, it.nx, it.nu, len(it.parameter_names), 0, it.N]

    class _Jac(casadi.Callback):
        def __init__(self):
            casadi.Callback.__init__(self)
            self.construct(f"jac_{name}_cb", {})

        def get_n_in(self):
            return 8  # the six inputs + the two nominal outputs

        def get_n_out(self):
            return 12  # jac_<out>_<in>, outputs outermost

        def get_sparsity_in(self, i):
            if i < 6:
                return casadi.Sparsity.dense(sizes_in[i], 1)
            return casadi.Sparsity.dense(it.nx, 1 if i == 6 else it.n_steps + 1)

        def get_sparsity_out(self, i):
            o, k = divmod(i, 6)
            rows = it.nx if o == 0 else it.nx * (it.n_steps + 1)
            if o == 0 and k in (1, 2, 3):
                return casadi.Sparsity.dense(rows, sizes_in[k])
            return casadi.Sparsity(rows, sizes_in[k])  # structurally zero: t_span, a, d, and every xall block

        def eval(self, arg):
            J = it.jacobian(np.array(arg[0]).ravel(), np.array(arg[1]).ravel(), np.array(arg[2]).ravel(),
                            np.array(arg[3]).ravel(), None, np.array(arg[5]).ravel())
            out = []
            for i in range(12):
                o, k = divmod(i, 6)
                blk = {1: J["x0"], 2: J["u"], 3: J["p"]}.get(k) if o == 0 else None
                sp = self.get_sparsity_out(i)
                out.append(casadi.DM(sp) if blk is None else casadi.DM(blk[..., 0].cpu().numpy()))
            return out

    class _Fun(casadi.Callback):
        def __init__(self):
            casadi.Callback.__init__(self)
            self._jac = None
            self.construct(f"{name}_cb", {})

        def get_n_in(self):
            return 6

        def get_n_out(self):
            return 2

        def get_name_in(self, i):
            return it.input_names[i]

        def get_name_out(self, i):
            return it.output_names[i]

        def get_sparsity_in(self, i):
            return casadi.Sparsity.dense(sizes_in[i], 1)

        def get_sparsity_out(self, i):
            return casadi.Sparsity.dense(it.nx, 1 if i == 0 else it.n_steps + 1)

        def eval(self, arg):
            r = it(np.array(arg[0]).ravel(), np.array(arg[1]).ravel(), np.array(arg[2]).ravel(),
                   np.array(arg[3]).ravel(), None, np.array(arg[5]).ravel())
            return [casadi.DM(r["xf"][:, 0].cpu().numpy()), casadi.DM(r["xall"][..., 0].cpu().numpy())]

        def has_jacobian(self):
            return True

        def get_jacobian(self, jname, inames, onames, opts):
            self._jac = _Jac()  # kept alive by the owner
            return self._jac

    return _Fun()


def GpuRK4(n_integration_steps=10, parameter_names=(), device=None):
    """``OdeSolver.RK4`` subclass instance whose integrator evaluates on the GPU; passes cocofest's
    ``isinstance(ode_solver, (OdeSolver.RK1, RK2, RK4, COLLOCATION))`` checks (fes_ocp.py:334-338)."""
    try:
        from bioptim import OdeSolver as _BioOdeSolver
    except ImportError as e:
        raise ImportError("GpuRK4 needs bioptim 3.2.1 (not installed); use FesShooting for batched defects") from e

    class _Integrator:
        def __init__(self, ode: dict, ode_opt: dict):
            self.model = ode_opt["model"]
            core = FesIntegrator(self.model, _Steps(4, n_integration_steps), parameter_names, device)
            self.core = core
            self.function = as_casadi_function(core)
            self.shape_xf, self.shape_xall = core.shape_xf, core.shape_xall
            self.step_times_from_dt = core.step_times_from_dt

        def __call__(self, *args, **kwargs):
            return self.function(*args, **kwargs)

        def map(self, *args, **kwargs):
            return self.function.map(*args, **kwargs)

    class _GpuRK4(_BioOdeSolver.RK4):
        @property
        def integrator(self):
            return _Integrator

    return _GpuRK4(n_integration_steps=n_integration_steps)


class _Steps:
    def __init__(self, scheme, n_integration_steps):
        self.scheme, self.n_integration_steps = scheme, n_integration_steps
