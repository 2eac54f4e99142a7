"""``IvpFes``: the reference's initial-value-problem driver (cocofest/integration/ivp_fes.py) on the GPU.

Same constructor dictionaries, defaults, checks and error messages; ``integrate()`` returns the reference's
``(states, time)`` pair -- ``states[key]`` of shape ``[1, ns*S+1]`` for key in Cn, F(, A, Tau1, Km) and the
matching time vector (bioptim Solution.integrate with SolutionMerge.NODES, duplicated_times=False).
Where the reference hacks a bioptim OptimalControlProgram to loop over nodes in Python
(ivp_fes.py:258-297 -> solve_ivp_interface.py:56-69), this class lowers the problem to one
``cocofest_b200.fes.ivp`` call.  ``integrate_batch`` is the data-parallel extension: many parameter sets /
initial states for the same protocol in one launch.
"""
from __future__ import annotations

import numpy as np
import torch

from . import fes as _fes
from .models import (DingModelPulseIntensityFrequency, DingModelPulseWidthFrequency, FesModel)
from .ode_solver import OdeSolver, scheme_and_steps


class IvpFes:
    def __init__(self, fes_parameters: dict = None, ivp_parameters: dict = None, device=None, dtype=torch.float64):
        self._fill_fes_dict(fes_parameters)
        self._fill_ivp_dict(ivp_parameters)
        self.dictionaries_check()

        self.model = self.fes_parameters["model"]
        self.stim_time = self.model.stim_time
        self.n_stim = len(self.stim_time)
        self.pulse_width = self.fes_parameters["pulse_width"]
        self.pulse_intensity = self.fes_parameters["pulse_intensity"]
        self.final_time = self.ivp_parameters["final_time"]
        self.n_shooting = _fes.prepare_n_shooting(self.stim_time, self.final_time)
        self.pulse_mode = self.fes_parameters["pulse_mode"]
        self._pulse_mode_settings()
        self.dt = np.array([self.final_time / self.n_shooting])

        self.controls_keys = None
        if isinstance(self.model, DingModelPulseWidthFrequency):
            self.controls_keys = ["last_pulse_width"]
        if isinstance(self.model, DingModelPulseIntensityFrequency):
            self.controls_keys = ["pulse_intensity"]

        data, stim_idx_at_node_list = self.model.get_numerical_data_time_series(self.n_shooting, self.final_time)
        self.numerical_data_time_series = data
        self.stim_idx_at_node_list = stim_idx_at_node_list
        self.window = np.ascontiguousarray(data["stim_time"][:, 0, :].T)  # [ns+1, N]
        self.x_init, self.u_init = self.build_initial_guess_from_ocp(self, stim_idx_at_node_list)

        self.ode_solver = self.ivp_parameters["ode_solver"]
        self.scheme, self.n_integration_steps = scheme_and_steps(self.ode_solver)
        self.n_threads = self.ivp_parameters["n_threads"]
        self.device = torch.device(device) if device is not None else torch.device("cuda")
        self.dtype = dtype

    # ---- dictionaries (ivp_fes.py:113-230) ---------------------------------------------------------------
    def _fill_fes_dict(self, fes_parameters):
        default = {"model": FesModel, "stim_time": None, "pulse_width": 0.0003, "pulse_intensity": 50, "pulse_mode": "single"}
        fes_parameters = {} if fes_parameters is None else fes_parameters
        if isinstance(fes_parameters, dict):
            for key, val in default.items():
                fes_parameters.setdefault(key, val)
        self.fes_parameters = fes_parameters

    def _fill_ivp_dict(self, ivp_parameters):
        default = {"final_time": None, "ode_solver": OdeSolver.RK4(n_integration_steps=10), "n_threads": 1}
        ivp_parameters = {} if ivp_parameters is None else ivp_parameters
        if isinstance(ivp_parameters, dict):
            for key, val in default.items():
                ivp_parameters.setdefault(key, val)
        self.ivp_parameters = ivp_parameters

    def dictionaries_check(self):
        fp, ip = self.fes_parameters, self.ivp_parameters
        if not isinstance(fp, dict):
            raise ValueError("fes_parameters must be a dictionary")
        if not isinstance(ip, dict):
            raise ValueError("ivp_parameters must be a dictionary")
        model = fp["model"]
        if not isinstance(model, FesModel):
            raise TypeError("model must be a FesModel type")

        def number_or_list(v):
            return isinstance(v, (int, float, list)) and not isinstance(v, bool)

        if isinstance(model, DingModelPulseWidthFrequency):
            pw = fp["pulse_width"]
            if not number_or_list(pw):
                raise TypeError("pulse_width must be int, float or list type")
            ok = all(w >= model.pd0 for w in pw) if isinstance(pw, list) else pw >= model.pd0
            if not ok:
                raise ValueError("pulse width must be greater than minimum pulse width")
        if isinstance(model, DingModelPulseIntensityFrequency):
            pi = fp["pulse_intensity"]
            if not number_or_list(pi):
                raise TypeError("pulse_intensity must be int, float or list type")
            lo = model.min_pulse_intensity()
            ok = all(i >= lo for i in pi) if isinstance(pi, list) else bool(pi >= lo)
            if not ok:
                raise ValueError("Pulse intensity must be greater than minimum pulse intensity")
        if not isinstance(fp["pulse_mode"], str):
            raise ValueError("pulse_mode must be a string type")
        if not isinstance(ip["final_time"], (int, float)) or isinstance(ip["final_time"], bool):
            raise ValueError("final_time must be an int or float type")
        try:
            scheme_and_steps(ip["ode_solver"])
        except (ValueError, TypeError):
            raise ValueError("ode_solver must be a OdeSolver type")
        if not isinstance(ip["n_threads"], int):
            raise ValueError("n_thread must be a int type")

    def _pulse_mode_settings(self):
        """ivp_fes.py:232-256 -- like the reference this mutates ``model.stim_time``."""
        if self.pulse_mode == "single":
            return
        self.stim_time = _fes.pulse_mode_times(self.stim_time, self.pulse_mode)
        self.model.stim_time = self.stim_time
        self.n_stim = len(self.stim_time)
        self.n_shooting = _fes.prepare_n_shooting(self.stim_time, self.final_time)

    # ---- controls per node (ivp_fes.py:312-357) ----------------------------------------------------------
    def build_initial_guess_from_ocp(self, ocp, stim_idx_at_node_list=None):
        """Returns (x_init [nx], u_init [ns, nu] | None)."""
        x = np.asarray(ocp.model.standard_rest_values(), dtype=np.float64).ravel()
        u = None
        ns, idx = self.n_shooting, stim_idx_at_node_list
        if ocp.controls_keys and "pulse_intensity" in ocp.controls_keys:
            N = self.model.sum_stim_truncation
            pi = self.pulse_intensity
            if isinstance(pi, list) and len(pi) != 1:
                u = np.empty((ns, N))
                for j in range(N):          # window slot j, right-aligned with the node's window
                    for i in range(ns):
                        src = idx[i][0] if i < ns - j else idx[i][(i - (ns - j)) + 1]
                        u[i, j] = pi[src]
            elif isinstance(pi, list):
                u = np.full((ns, N), float(pi[0]))
            else:
                # the reference reads self.pulse_width here (ivp_fes.py:339), a bug we do not replicate
                u = np.full((ns, N), float(pi))
        if ocp.controls_keys and "last_pulse_width" in ocp.controls_keys:
            pw = self.pulse_width
            if isinstance(pw, list) and len(pw) != 1:
                u = np.array([[pw[idx[i][-1]]] for i in range(len(idx) - 1)], dtype=np.float64)
            else:
                u = np.full((ns, 1), float(pw[0] if isinstance(pw, list) else pw))
        return x, u

    # ---- integration -------------------------------------------------------------------------------------
    def _time_vector(self):
        S, ns, dt = self.n_integration_steps, self.n_shooting, float(self.dt[0])
        t = np.concatenate([np.linspace(dt * n, dt * n + dt, S + 1)[:-1] for n in range(ns)] + [[dt * ns]])
        return t

    def integrate(self, shooting_type=None, integrator=None, to_merge=None, return_time=True, duplicated_times=False):
        """Single shooting from the rest state through all nodes; every RK sub-step is returned.

        The keyword arguments of the reference signature are accepted; only the default combination
        (Shooting.SINGLE, SolutionIntegrator.OCP, merge NODES+PHASES, duplicated_times=False) is implemented."""
        if duplicated_times:
            raise NotImplementedError("duplicated_times=True is not provided by the GPU path")
        res = _fes.ivp(self.model.kind, self.model.with_fatigue, self.model.param_vector(),
                       torch.as_tensor(self.x_init, device=self.device)[:, None], torch.as_tensor(self.window, device=self.device),
                       float(self.dt[0]), n_steps=self.n_integration_steps,
                       ctrl=None if self.u_init is None else torch.as_tensor(self.u_init, device=self.device),
                       n_shooting=self.n_shooting, scheme=self.scheme, tauc=self.model.tauc,
                       r0_km=self.model.r0_km_relationship, trace="steps", dtype=self.dtype)
        xall = res.trace[:, :, 0].T.cpu().numpy()  # [nx, npts]
        states = {name: xall[i:i + 1, :] for i, name in enumerate(self.model.name_dof)}
        if return_time:
            return states, self._time_vector()
        return states

    def integrate_batch(self, params=None, x0=None, dirs=0, trace="nodes", trace_batch=None, sens_trace=False):
        """Data-parallel extension: integrate the protocol of this problem for B parameter sets / initial states.

        params: None | tensor [14, B] (canonical order, ``cocofest_b200.fes.PARAM_NAMES``);  x0: None | [nx, B].
        Returns ``cocofest_b200.fes.IvpResult`` (tensors on ``self.device``)."""
        if params is None and x0 is None:
            raise ValueError("give params [14, B] and / or x0 [nx, B]")
        B = (params if params is not None else x0).shape[-1]
        if params is None:
            params = torch.as_tensor(self.model.param_vector())
        if x0 is None:
            p = torch.as_tensor(params, dtype=torch.float64)
            p = p if p.dim() == 2 else p[:, None].expand(-1, B)
            x0 = torch.zeros((self.model.nb_state, B), dtype=torch.float64, device=p.device)
            if self.model.with_fatigue:
                x0[2], x0[3], x0[4] = p[0], p[2], p[1]  # A_rest (a_scale), tau1_rest, km_rest
        return _fes.ivp(self.model.kind, self.model.with_fatigue, torch.as_tensor(params).to(self.device),
                        torch.as_tensor(x0).to(self.device), torch.as_tensor(self.window, device=self.device),
                        float(self.dt[0]), n_steps=self.n_integration_steps,
                        ctrl=None if self.u_init is None else torch.as_tensor(self.u_init, device=self.device),
                        n_shooting=self.n_shooting, scheme=self.scheme, dirs=dirs, tauc=self.model.tauc,
                        r0_km=self.model.r0_km_relationship, trace=trace, trace_batch=trace_batch,
                        sens_trace=sens_trace, dtype=self.dtype)

    # ---- alternate constructors (ivp_fes.py:359-443) -------------------------------------------------------
    @classmethod
    def from_frequency_and_final_time(cls, fes_parameters: dict = None, ivp_parameters: dict = None, **kw):
        frequency = fes_parameters["frequency"]
        if not isinstance(frequency, int):
            raise ValueError("Frequency must be an int")
        round_down = fes_parameters["round_down"]
        if not isinstance(round_down, bool):
            raise ValueError("Round down must be a bool")
        final_time = ivp_parameters["final_time"]
        if not isinstance(final_time, (int, float)):
            raise ValueError("Final time must be an int or float")
        n_stim = final_time * frequency
        if round_down or float(n_stim).is_integer():
            n_stim = int(n_stim)
        else:
            raise ValueError(
                "The number of stimulation needs to be integer within the final time t, set round down "
                "to True or set final_time * frequency to make the result an integer.")
        fes_parameters["n_stim"] = n_stim
        fes_parameters["stim_time"] = list(np.round([i / frequency for i in range(n_stim)], 3))
        return cls(fes_parameters, ivp_parameters, **kw)

    @classmethod
    def from_frequency_and_n_stim(cls, fes_parameters: dict = None, ivp_parameters: dict = None, **kw):
        n_stim = fes_parameters["n_stim"]
        if not isinstance(n_stim, int):
            raise ValueError("n_stim must be an int")
        frequency = fes_parameters["frequency"]
        if not isinstance(frequency, int):
            raise ValueError("Frequency must be an int")
        ivp_parameters["final_time"] = n_stim / frequency
        fes_parameters["stim_time"] = list(np.round([i / frequency for i in range(n_stim)], 3))
        return cls(fes_parameters, ivp_parameters, **kw)
