"""Force-parameter identification on the GPU behind the reference's class surface.

Reference surface mirrored (cocofest tree):
  identification/identification_method.py:5-78 (full_data_extraction), :80-170 (average_data_extraction),
  :173-189 (sparse_data_extraction raises NotImplementedError), :253-278 (force_at_node_in_ocp);
  identification/ding2003_force_parameter_identification.py:25-427, ding2007_...:25-345, hmed2018_...:25-365
  (constructor arguments, default boxes / initial guesses, input checks, ``force_model_identification()``);
  on-disk format: pickles ``{"time", "force", "stim_time"}`` + ``"pulse_width"`` | ``"pulse_intensity"``
  (identification_method.py:26-41, examples/identification/*.py:43-47).

What differs, on purpose: the reference builds ``OcpFesId`` (multiple shooting, states as decision variables)
and runs ONE Ipopt solve; here the same objective -- sum over all nodes of dt (F(t_j) - F_data(t_j))^2
(fes_identification_ocp.py:248-266) -- is minimised over the same box by the on-device bounded
Levenberg-Marquardt iteration of ``fes_identify_*`` from ``n_starts`` starting points at once (the reference's
initial guess first, the others drawn log-uniformly inside the box), and the best candidate is returned.
Both reach the same minimiser when the data are informative; no Ipopt, casadi or bioptim is involved.

Deliberate deviations from the reference on this path, all three pinned by ``tests/test_parity_deviations.py`` and
listed in DESIGN.md section 5:
  1. Hmed2018 per-node intensity matrix: slot ``j`` of node ``n`` holds the intensity of the pulse that sits in slot ``j``
     of that node's window (right-aligned, as the window itself).  The reference's ``OcpFesId._set_u_bounds``
     (fes_identification_ocp.py:332-343) indexes ``control_value[stim_idx[j-i][j-i]]`` / ``control_value[stim_idx[i][0]]``,
     which pairs pulses with other pulses' intensities as soon as the intensities differ (and runs off the index lists
     for most (i, j)); identified ``ar / bs / Is / cr`` therefore differ from the reference's on multi-intensity data.
  2. Node assignment of pulses (``fes.stim_windows``): ``t_i <= n dt`` up to ``1e-9 dt`` instead of the reference's raw
     float compare (ding2003.py:416), which drops the pulse at 0.1 from node 1 when dt = 0.3 / 3; the reference's own
     golden vectors correspond to the grid-exact assignment.
  3. ``IvpFes`` with a scalar ``pulse_intensity`` fills the control matrix with that scalar; the reference reads
     ``self.pulse_width`` there (ivp_fes.py:339).
"""
from __future__ import annotations

import math
import pickle
from dataclasses import dataclass

import numpy as np
import torch

from . import fes as _fes
from ._lib import FES_DING2003, FES_DING2007, FES_HMED2018, PARAM_INDEX
from .models import (DingModelFrequency, DingModelPulseIntensityFrequency, DingModelPulseWidthFrequency, FesModel)
from .ode_solver import OdeSolver, scheme_and_steps


# ---- data extraction (identification_method.py) ------------------------------------------------------------
def _load(path):
    with open(path, "rb") as f:
        return pickle.load(f)


def full_data_extraction(model_data_path):
    """Concatenates the trains of several pickles on one time axis (identification_method.py:5-78).
    Returns (time, stim_time, force, discontinuity_phase_list) as flat lists."""
    g_force, g_stim, g_time, discontinuity = [], [], [], []
    for i, path in enumerate(model_data_path):
        data = _load(path)
        t0 = data["stim_time"][0]
        stim = list(data["stim_time"]) if t0 == 0 else [s - t0 for s in data["stim_time"]]
        # the reference treats "time" as flat when t0 == 0 and as a list of rows otherwise (:37-41); both layouts are
        # accepted here and flattened per file, which is what its final list expansion (:71-73) assumes
        time = [float(t) - (t0 if t0 != 0 else 0.0) for t in np.ravel(np.asarray(data["time"], dtype=np.float64))]
        if i != 0:
            discontinuity.append(len(g_stim[-1]) if not discontinuity else discontinuity[-1] + len(g_stim[-1]))
            # the reference shifts the stimulation times twice (identification_method.py:52-59); kept as is
            last = g_time[i - 1][-1]
            stim = [s + last for s in stim]
            time = [(t + last) for t in time]
            stim = [s + last for s in stim]
        g_force.append([float(x) for x in np.ravel(np.asarray(data["force"], dtype=np.float64))])
        g_stim.append(stim)
        g_time.append(time)
    flat = lambda rows: [x for row in rows for x in row]
    return flat(g_time), flat(g_stim), flat(g_force), discontinuity


def average_data_extraction(model_data_path, train_duration=1):
    """Averages the force curves of each pickle (identification_method.py:80-170).  ``train_duration`` is a free
    name in the reference's body (NameError as written); it is an argument here, 1 s like ``train_width``."""
    g_force, g_stim, g_time, discontinuity = [], [], [], []
    for i, path in enumerate(model_data_path):
        data = _load(path)
        force = data["force"]
        st = data["stim_time"]
        threshold = st[1] - st[0]
        gaps = [st[j] - st[j - 1] for j in range(1, len(st)) if st[j] - st[j - 1] < threshold * 1.5]
        frequency = round(1 / np.mean(gaps), 0)
        time = data["time"] if st[0] == 0 else [[(t - st[0]) for t in row] for row in data["time"]]
        smallest = min(len(row) for row in force)
        force = np.mean([row[:smallest] for row in force], axis=0).tolist()
        time = [x for row in time for x in row][:smallest]
        stim = list(np.linspace(0, 1, int(frequency * train_duration) + 1)[:-1])
        if i == len(model_data_path) - 1:
            stim = stim + [time[-1]]
        if i != 0:
            discontinuity.append(len(g_stim[-1]) if not discontinuity else discontinuity[-1] + len(g_stim[-1]))
            last = g_time[i - 1][-1]
            time = [(t + last) for t in time]
            stim = [(s + last) for s in stim]
        g_force.append(force)
        g_stim.append(stim)
        g_time.append(time)
    flat = lambda rows: [x for row in rows for x in row]
    return flat(g_time), flat(g_stim), flat(g_force), discontinuity


def sparse_data_extraction(model_data_path, force_curve_number=5):
    """identification_method.py:173-189."""
    raise NotImplementedError("This method has not been tested yet")


def force_at_node_in_ocp(time, force, n_shooting, final_time, sparse=None):
    """Force data resampled at the ns+1 nodes (identification_method.py:253-278)."""
    temp_time = np.linspace(0, final_time, n_shooting + 1)
    return list(np.interp(temp_time, time, force))


@dataclass
class IdentificationResult:
    """What ``force_identification_result`` exposes here (the reference stores a bioptim Solution)."""
    parameters: dict          # key -> identified value (best candidate)
    cost: float               # sum_j dt (F(t_j) - F_data(t_j))^2 at the optimum
    time: np.ndarray          # node times [ns+1]
    force: np.ndarray         # model force at the nodes for the identified parameters
    force_data: np.ndarray    # the tracked data at the nodes
    all_params: torch.Tensor  # [14, n_starts] every candidate's end point
    all_costs: torch.Tensor   # [n_starts]
    evaluations: torch.Tensor  # [n_starts] cost evaluations spent per candidate
    hessian: np.ndarray       # Gauss-Newton matrix 2 J^T W J of the best candidate (P-column order)
    columns: tuple


# ---- the identification classes ------------------------------------------------------------------------------
class ParameterIdentification:
    """Shared machinery of the three reference classes (identification_abstract_class.py + the concrete files)."""

    model_type = DingModelFrequency
    kind = FES_DING2003
    DEFAULTS = {  # key: (initial_guess, min_bound, max_bound)   ding2003_force_parameter_identification.py:121-150
        "a_rest": (1000, 1, 10000), "km_rest": (0.5, 0.001, 1), "tau1_rest": (0.5, 0.0001, 1), "tau2": (0.5, 0.0001, 1),
    }

    def __init__(self, model, data_path=None, identification_method: str = "full", double_step_identification: bool = False,
                 key_parameter_to_identify: list = None, additional_key_settings: dict = None, final_time: float = 1,
                 objective: dict = None, use_sx: bool = True, ode_solver=None, n_threads: int = 1, control_type=None,
                 n_starts: int = 256, max_iter: int = 200, seed: int = 0, device=None, **kwargs):
        self.default_values = self._set_default_values(model)
        additional_key_settings = {} if additional_key_settings is None else additional_key_settings
        self.input_sanity(model, data_path, identification_method, double_step_identification,
                          key_parameter_to_identify, additional_key_settings)
        # identified parameters are blanked on the model like the reference does (ding2003_...:77-83)
        for key in key_parameter_to_identify:
            setattr(model, key, None)
        self.model = model
        self.key_parameter_to_identify = key_parameter_to_identify
        self.additional_key_settings = self.key_setting_to_dictionary(additional_key_settings)
        self.data_path = [data_path] if isinstance(data_path, str) else data_path
        self.force_model_identification_method = identification_method
        self.double_step_identification = double_step_identification
        self.final_time = final_time
        self.objective = objective
        self.ode_solver = OdeSolver.RK4(n_integration_steps=10) if ode_solver is None else ode_solver
        self.use_sx, self.n_threads, self.control_type, self.kwargs = use_sx, n_threads, control_type, kwargs
        self.n_starts, self.max_iter, self.seed = int(n_starts), int(max_iter), int(seed)
        self.device = torch.device(device) if device is not None else torch.device("cuda")
        self.force_ocp = None  # reference attribute; no OCP object exists on this path
        self.force_identification_result = None

    # -- defaults / checks (ding2003_...:107-275) --------------------------------------------------------------
    def _set_default_values(self, model):
        return {k: {"initial_guess": g, "min_bound": lo, "max_bound": hi, "function": getattr(model, "set_" + k, None), "scaling": 1}
                for k, (g, lo, hi) in self.DEFAULTS.items()}

    def input_sanity(self, model=None, data_path=None, identification_method=None, double_step_identification=None,
                     key_parameter_to_identify=None, additional_key_settings=None):
        if not isinstance(model, FesModel) or model.kind != self.kind:
            raise TypeError(f"The given model must be a {self.model_type.__name__}")
        if model._with_fatigue:
            raise ValueError("The given model is not valid and should not be including the fatigue equation in the model")
        self.check_experiment_force_format(data_path)
        if identification_method not in ["full", "average", "sparse"]:
            raise ValueError("The given model identification method is not valid,only 'full', 'average' and 'sparse' are available,"
                             f" the given value is {identification_method}")
        if not isinstance(double_step_identification, bool):
            raise TypeError("The given double_step_identification must be bool type,"
                            f" the given value is {type(double_step_identification)} type")
        if isinstance(key_parameter_to_identify, list):
            for key in key_parameter_to_identify:
                if key not in self.default_values:
                    raise ValueError(f"The given key_parameter_to_identify is not valid, the given value is {key},"
                                     f" the available values are {list(self.default_values.keys())}")
        else:
            raise TypeError("The given key_parameter_to_identify must be list type,"
                            f" the given value is {type(key_parameter_to_identify)} type")
        if isinstance(additional_key_settings, dict):
            for key, settings in additional_key_settings.items():
                if key not in self.default_values:
                    raise ValueError(f"The given additional_key_settings is not valid, the given value is {key},"
                                     f" the available values are {list(self.default_values.keys())}")
                for name, value in settings.items():
                    if name not in self.default_values[key]:
                        raise ValueError(f"The given additional_key_settings is not valid, the given value is {name},"
                                         f" the available values are {list(self.default_values[key].keys())}")
                    if not isinstance(value, type(self.default_values[key][name])):
                        raise TypeError(f"The given additional_key_settings value is not valid, the given value is {type(value)},"
                                        f" the available type is {type(self.default_values[key][name])}")
        else:
            raise TypeError("The given additional_key_settings must be dict type,"
                            f" the given value is {type(additional_key_settings)} type")

    def key_setting_to_dictionary(self, key_settings):
        out = {}
        for key in self.key_parameter_to_identify:
            out[key] = {name: (key_settings[key][name] if key in key_settings and name in key_settings[key] else default)
                        for name, default in self.default_values[key].items()}
            lo, hi = float(out[key]["min_bound"]), float(out[key]["max_bound"])
            if not (math.isfinite(lo) and math.isfinite(hi) and lo < hi):
                raise ValueError(f"{key}: min_bound must be finite and smaller than max_bound, got [{lo}, {hi}]")
            g = float(out[key]["initial_guess"])
            if not (lo <= g <= hi):
                raise ValueError(f"{key}: initial_guess {g} lies outside [{lo}, {hi}]")
        return out

    @staticmethod
    def check_experiment_force_format(data_path):
        if isinstance(data_path, list):
            for i, p in enumerate(data_path):
                if not isinstance(p, str):
                    raise TypeError(f"In the given list, all model_data_path must be str type, path index n°{i} is not str type")
                if not p.endswith(".pkl"):
                    raise TypeError("In the given list, all model_data_path must be pickle type and end with .pkl,"
                                    f" path index n°{i} is not ending with .pkl")
        elif isinstance(data_path, str):
            if not data_path.endswith(".pkl"):
                raise TypeError("In the given list, all model_data_path must be pickle type and end with .pkl,"
                                " path index is not ending with .pkl")
        else:
            raise TypeError(f"In the given path, model_data_path must be str or list[str] type, the input is {type(data_path)} type")

    # -- data -> node grid -------------------------------------------------------------------------------------------
    def _extract(self, method):
        if method == "full":
            return full_data_extraction(self.data_path)
        if method == "average":
            return average_data_extraction(self.data_path)
        return sparse_data_extraction(self.data_path, self.kwargs.get("force_curve_number", 5))

    def _control_source(self, method):
        """Per-pulse control values read from the pickles (None for ding2003)."""
        return None

    def _controls(self, values, stim_idx, ns, N):
        return None

    # -- the identification ------------------------------------------------------------------------------------------
    def _starting_points(self, base):
        """[14, n_starts]: the configured initial guess first, then log-uniform draws inside the box."""
        keys = self.key_parameter_to_identify
        P = torch.tensor(base, dtype=torch.float64)[:, None].repeat(1, self.n_starts)
        g = torch.Generator().manual_seed(self.seed)
        for key in keys:
            s = self.additional_key_settings[key]
            lo, hi = float(s["min_bound"]), float(s["max_bound"])
            row = PARAM_INDEX[key]
            r = torch.rand(self.n_starts, generator=g, dtype=torch.float64)
            if lo > 0.0:   # positive box: log-uniform, the boxes span up to five decades
                draws = torch.exp(math.log(lo) + (math.log(hi) - math.log(lo)) * r)
            else:          # a box that reaches zero or below has no logarithm: uniform
                draws = lo + (hi - lo) * r
            draws[0] = float(s["initial_guess"])
            P[row] = draws
        return P

    def _identify(self, method, initial_guess=None):
        time, stim, force, _ = self._extract(method)
        stim = list(np.round(stim, 3))  # ding2003_...:403
        ns = _fes.prepare_n_shooting(stim, self.final_time)
        target = np.asarray(force_at_node_in_ocp(time, force, ns, self.final_time), dtype=np.float64)
        N = self.model.sum_stim_truncation
        self.model.stim_time = stim
        data, stim_idx = self.model.get_numerical_data_time_series(ns, self.final_time)
        win = np.ascontiguousarray(data["stim_time"][:, 0, :].T)
        ctrl = self._controls(self._control_source(method), stim_idx, ns, N)
        scheme, n_steps = scheme_and_steps(self.ode_solver)
        base = _fes.default_params(self.kind)
        for key, idx in PARAM_INDEX.items():  # fixed (not identified) parameters come from the model
            v = getattr(self.model, key, None)
            if v is not None and not (key == "a_rest" and self.kind == FES_DING2007):
                base[idx] = float(v)
        if initial_guess:
            for key, v in initial_guess.items():
                self.additional_key_settings[key]["initial_guess"] = float(v)
        starts = self._starting_points(base).to(self.device)
        lower = {k: float(self.additional_key_settings[k]["min_bound"]) for k in self.key_parameter_to_identify}
        upper = {k: float(self.additional_key_settings[k]["max_bound"]) for k in self.key_parameter_to_identify}
        res = _fes.identify(self.kind, starts, torch.as_tensor(win, device=self.device), self.final_time / ns,
                            torch.as_tensor(target, device=self.device), n_steps=n_steps,
                            ctrl=None if ctrl is None else torch.as_tensor(ctrl, device=self.device),
                            identify=self.key_parameter_to_identify, lower=lower, upper=upper, max_iter=self.max_iter,
                            scheme=scheme, tauc=self.model.tauc, r0_km=self.model.r0_km_relationship)
        b = res.best()
        best = res.params[:, b].contiguous()
        traj = _fes.ivp(self.kind, False, best, torch.zeros((2, 1), dtype=torch.float64, device=self.device),
                        torch.as_tensor(win, device=self.device), self.final_time / ns, n_steps=n_steps,
                        ctrl=None if ctrl is None else torch.as_tensor(ctrl, device=self.device), n_shooting=ns,
                        scheme=scheme, tauc=self.model.tauc, r0_km=self.model.r0_km_relationship, trace="nodes")
        params = {k: float(best[PARAM_INDEX[k]]) for k in self.key_parameter_to_identify}
        self.force_identification_result = IdentificationResult(
            parameters=params, cost=float(res.cost[b]), time=np.linspace(0, self.final_time, ns + 1),
            force=traj.trace[:, 1, 0].cpu().numpy(), force_data=target, all_params=res.params, all_costs=res.cost,
            evaluations=res.evaluations, hessian=res.hessian()[:, :, b].cpu().numpy(), columns=res.columns)
        return params

    def _force_model_identification_for_initial_guess(self):
        """First step of the double-step identification: on the averaged data (ding2003_...:333-377)."""
        return self._identify("average")

    def force_model_identification(self):
        """Returns {key: identified value} and stores the details in ``force_identification_result``."""
        guess = self._force_model_identification_for_initial_guess() if self.double_step_identification else None
        params = self._identify(self.force_model_identification_method, guess)
        for key, v in params.items():  # the reference's parameter functions write the values back into the model
            setattr(self.model, key, v)
        return params


class DingModelFrequencyForceParameterIdentification(ParameterIdentification):
    """identification/ding2003_force_parameter_identification.py:25."""


class DingModelPulseWidthFrequencyForceParameterIdentification(ParameterIdentification):
    """identification/ding2007_force_parameter_identification.py:25; pickles carry ``"pulse_width"``."""

    model_type = DingModelPulseWidthFrequency
    kind = FES_DING2007
    DEFAULTS = {  # ding2007_force_parameter_identification.py:105-148
        "tau1_rest": (0.5, 0.0001, 1), "tau2": (0.5, 0.0001, 1), "km_rest": (0.5, 0.001, 1), "a_scale": (5000, 1, 10000),
        "pd0": (1e-4, 1e-4, 6e-4), "pdt": (1e-4, 1e-4, 6e-4),
    }

    @staticmethod
    def pulse_width_extraction(data_path) -> list:
        """ding2007_...:190-212."""
        return [w for p in data_path for w in _load(p)["pulse_width"]]

    def _control_source(self, method):
        pw = self.pulse_width_extraction(self.data_path)
        return float(np.mean(np.array(pw))) if method == "average" else pw  # ding2007_...:300-309

    def _controls(self, values, stim_idx, ns, N):
        if isinstance(values, list) and len(values) != 1:  # width of the LAST pulse of each node's window
            return np.array([[values[stim_idx[n][-1]]] for n in range(ns)], dtype=np.float64)
        return np.full((ns, 1), float(values[0] if isinstance(values, list) else values))


class DingModelPulseIntensityFrequencyForceParameterIdentification(ParameterIdentification):
    """identification/hmed2018_force_parameter_identification.py:25; pickles carry ``"pulse_intensity"``."""

    model_type = DingModelPulseIntensityFrequency
    kind = FES_HMED2018
    DEFAULTS = {  # hmed2018_force_parameter_identification.py:103-160
        "a_rest": (1000, 1, 10000), "km_rest": (0.5, 0.001, 1), "tau1_rest": (0.5, 0.0001, 1), "tau2": (0.5, 0.0001, 1),
        "ar": (0.5, 0.01, 1), "bs": (0.05, 0.001, 0.1), "Is": (50, 1, 150), "cr": (1, 0.01, 2),
    }

    @staticmethod
    def pulse_intensity_extraction(data_path) -> list:
        """hmed2018_...:215-237."""
        return [i for p in data_path for i in _load(p)["pulse_intensity"]]

    def _control_source(self, method):
        pi = self.pulse_intensity_extraction(self.data_path)
        return float(np.mean(np.array(pi))) if method == "average" else pi  # hmed2018_...:322-331

    def _controls(self, values, stim_idx, ns, N):
        if not (isinstance(values, list) and len(values) != 1):
            return np.full((ns, N), float(values[0] if isinstance(values, list) else values))
        # slot j of node n holds the intensity of the pulse in that slot of the node's window (right-aligned);
        # slots older than the first pulse keep the first pulse's intensity, irrelevant behind the -1e7 pads
        u = np.empty((ns, N))
        for n in range(ns):
            idx = stim_idx[n]
            u[n, :] = values[idx[0]]
            for s, src in enumerate(idx):
                u[n, N - len(idx) + s] = values[src]
        return u
