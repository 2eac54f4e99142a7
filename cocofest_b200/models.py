"""Host-side mirror of cocofest's six FES model classes (names, attributes, setters, rest values,
window builder) without casadi / bioptim.

Reference surface mirrored (cocofest tree): models/ding2003.py:17-150,394-429, ding2003_with_fatigue.py:17-136,
ding2007.py:17-119, ding2007_with_fatigue.py:16-130, hmed2018.py:17-118,303-316, hmed2018_with_fatigue.py:16-122,
models/model_maker.py:9-22, models/fes_model.py:8.

In the reference these classes build casadi expressions; here they are parameter containers that describe a
model to the CUDA integrator (``kind``, ``with_fatigue``, ``param_vector()``, ``get_numerical_data_time_series``).
The scalar helpers (``exp_time_fun`` ... ``system_dynamics``) evaluate one point with Python floats, as the
reference's ``tests/shard1/test_models_dynamics_without_bioptim.py`` uses them; they are not on the hot path --
every integration goes through ``cocofest_b200.fes`` to the GPU.
"""
from __future__ import annotations

import math
from abc import ABC, abstractmethod

import numpy as np

from . import fes as _fes
from ._lib import FES_DING2003, FES_DING2007, FES_HMED2018, PARAM_INDEX


class FesModel(ABC):
    """Abstract base (reference: models/fes_model.py:8)."""

    kind: int = None

    def __init__(self):
        self.stim_time = None

    @abstractmethod
    def standard_rest_values(self):
        ...

    @abstractmethod
    def serialize(self):
        ...


class DingModelFrequency(FesModel):
    """Ding 2003, stimulation frequency only (reference: models/ding2003.py:17)."""

    kind = FES_DING2003
    _fatigue = False
    _default_name = "ding2003"

    def __init__(self, model_name: str = None, muscle_name: str = None, stim_time: list = None,
                 previous_stim: dict = None, sum_stim_truncation: int = 20):
        super().__init__()
        self._model_name = self._default_name if model_name is None else model_name
        self._muscle_name = muscle_name
        self._sum_stim_truncation = sum_stim_truncation
        self._with_fatigue = self._fatigue
        self.pulse_apparition_time = None
        self.stim_time = list(stim_time) if stim_time else []
        self.previous_stim = previous_stim if previous_stim else {"time": []}
        self.all_stim = list(self.previous_stim["time"]) + self.stim_time
        d = _fes.default_params(FES_DING2003)
        self.tauc = _fes.DEFAULT_TAUC[FES_DING2003]
        self.r0_km_relationship = _fes.R0_KM_RELATIONSHIP
        self.a_rest = float(d[PARAM_INDEX["a_rest"]])
        self.tau1_rest = float(d[PARAM_INDEX["tau1_rest"]])
        self.tau2 = float(d[PARAM_INDEX["tau2"]])
        self.km_rest = float(d[PARAM_INDEX["km_rest"]])

    # -- setters keep bioptim's (model, value) signature (fes_identification_ocp.py:269-299 passes them) --
    def set_a_rest(self, model, a_rest):
        self.a_rest = a_rest

    def set_km_rest(self, model, km_rest):
        self.km_rest = km_rest

    def set_tau1_rest(self, model, tau1_rest):
        self.tau1_rest = tau1_rest

    def set_tau2(self, model, tau2):
        self.tau2 = tau2

    # -- descriptors --------------------------------------------------------------------------------
    def _dof(self, names, with_muscle_name=False):
        suffix = "_" + self.muscle_name if self.muscle_name and with_muscle_name else ""
        return [n + suffix for n in names]

    @property
    def name_dof(self) -> list:
        return self._dof(["Cn", "F", "A", "Tau1", "Km"][: self.nb_state])

    @property
    def nb_state(self) -> int:
        return 5 if self._with_fatigue else 2

    @property
    def model_name(self):
        return self._model_name

    @property
    def muscle_name(self):
        return self._muscle_name

    @property
    def with_fatigue(self) -> bool:
        return self._with_fatigue

    @property
    def sum_stim_truncation(self):
        return self._sum_stim_truncation

    _ID_KEYS = ("a_rest", "tau1_rest", "km_rest", "tau2")
    _FATIGUE_KEYS = ("alpha_a", "alpha_tau1", "alpha_km", "tau_fat")

    @property
    def identifiable_parameters(self) -> dict:
        keys = self._ID_KEYS + (self._FATIGUE_KEYS if self._with_fatigue else ())
        return {k: getattr(self, k) for k in keys}

    @property
    def km_name(self) -> str:
        return "Km" + ("_" + self.muscle_name if self.muscle_name else "")

    @property
    def cn_sum_name(self) -> str:
        return "Cn_sum" + ("_" + self.muscle_name if self.muscle_name else "")

    def _serial_keys(self):
        return ("tauc", "a_rest", "tau1_rest", "km_rest", "tau2") + (self._FATIGUE_KEYS if self._with_fatigue else ())

    def serialize(self):
        return type(self), {k: getattr(self, k) for k in self._serial_keys()}

    def standard_rest_values(self) -> np.ndarray:
        if self._with_fatigue:
            return np.array([[0], [0], [self._a_for_rest()], [self.tau1_rest], [self.km_rest]])
        return np.array([[0], [0]])

    def _a_for_rest(self):
        return self.a_rest

    # -- what the CUDA integrator needs --------------------------------------------------------------
    def param_vector(self) -> np.ndarray:
        """Canonical 14-vector of include/fes_b200.h from the model attributes."""
        p = _fes.default_params(self.kind)
        for key, idx in PARAM_INDEX.items():
            if key == "a_scale" or (key == "a_rest" and self.kind == FES_DING2007):
                continue
            if hasattr(self, key):
                p[idx] = float(getattr(self, key))
        if self.kind == FES_DING2007:
            p[PARAM_INDEX["a_rest"]] = float(self.a_scale)
        return p

    def _get_additional_previous_stim_time(self):
        while len(self.previous_stim["time"]) < self._sum_stim_truncation:
            self.previous_stim["time"].insert(0, _fes.PAD_TIME)
        return self.previous_stim

    def get_numerical_data_time_series(self, n_shooting, final_time, all_stim_time=None):
        """Per-node stimulation window, reference layout ``{"stim_time": [N, 1, ns+1]}`` plus
        ``stim_idx_at_node_list`` (ding2003.py:399-429)."""
        if self._sum_stim_truncation is None:
            raise ValueError("sum_stim_truncation must be given for this model (ding2007.py:36 defaults to None)")
        self.previous_stim = self._get_additional_previous_stim_time()
        stim_time = list(all_stim_time) if all_stim_time else list(self.stim_time)
        self.all_stim = list(self.previous_stim["time"]) + stim_time
        # previous_stim is already padded: pass it whole so the window builder adds nothing
        win, stim_idx = _fes.stim_windows(stim_time, n_shooting, final_time, self._sum_stim_truncation,
                                          previous_time=self.previous_stim["time"])
        return {"stim_time": np.transpose(win[:, np.newaxis, :], (2, 1, 0)).copy()}, stim_idx

    # -- scalar helpers (one point, Python floats) ---------------------------------------------------
    def get_r0(self, km):
        return km + self.r0_km_relationship

    @staticmethod
    def get_lambda_i(nb_stim: int, pulse_intensity=None) -> list:
        return [1 for _ in range(nb_stim)]

    def exp_time_fun(self, t, t_stim_i):
        return math.exp(-(t - t_stim_i) / self.tauc)

    def ri_fun(self, r0, time_between_stim):
        return 1 + (r0 - 1) * math.exp(-time_between_stim / self.tauc)

    def cn_sum_fun(self, r0, t, t_stim_prev, lambda_i=None):
        n = len(t_stim_prev)
        lam = [1] * n if lambda_i is None else lambda_i
        total = 0.0
        for i in range(n):
            ri = 1 if i == 0 else self.ri_fun(r0, t_stim_prev[i] - t_stim_prev[i - 1])
            total += ri * self.exp_time_fun(t, t_stim_prev[i]) * lam[i]
        return total

    def cn_dot_fun(self, cn, cn_sum):
        return (1 / self.tauc) * cn_sum - (cn / self.tauc)

    def calculate_cn_dot(self, cn, t, t_stim_prev, pulse_intensity=None):
        r0 = self.get_r0(self.km_rest)
        lam = self.get_lambda_i(len(t_stim_prev), pulse_intensity)
        return self.cn_dot_fun(cn, self.cn_sum_fun(r0, t, t_stim_prev, lam))

    def f_dot_fun(self, cn, f, a, tau1, km, force_length_relationship=1, force_velocity_relationship=1,
                  passive_force_relationship=0):
        q = cn / (km + cn)
        return (a * q - f / (tau1 + self.tau2 * q)) * (
            force_length_relationship * force_velocity_relationship + passive_force_relationship)

    def _a_input(self, a, **kw):
        return a

    def system_dynamics(self, cn, f, a=None, tau1=None, km=None, t=None, t_stim_prev=None, pulse_width=None,
                        pulse_intensity=None, force_length_relationship=1, force_velocity_relationship=1,
                        passive_force_relationship=0) -> np.ndarray:
        """dx/dt at one point (reference: system_dynamics of each model)."""
        cn_dot = self.calculate_cn_dot(cn, t, t_stim_prev, pulse_intensity)
        if self._with_fatigue:
            a_s, tau1_s, km_s = a, tau1, km
        else:
            a_s, tau1_s, km_s = self._a_for_rest(), self.tau1_rest, self.km_rest
        a_eff = self._a_input(a_s, pulse_width=pulse_width)
        f_dot = self.f_dot_fun(cn, f, a_eff, tau1_s, km_s, force_length_relationship, force_velocity_relationship,
                               passive_force_relationship)
        if not self._with_fatigue:
            return np.array([cn_dot, f_dot])
        return np.array([cn_dot, f_dot, self.a_dot_fun(a, f), self.tau1_dot_fun(tau1, f), self.km_dot_fun(km, f)])


    # -- bioptim hooks (ding2003.py:313-392 and the same block of every model) -----------------------------------------
    @staticmethod
    def dynamics(time, states, controls, parameters, algebraic_states, numerical_timeseries, nlp=None, fes_model=None,
                 force_length_relationship=1, force_velocity_relationship=1, passive_force_relationship=0):
        """The reference's ``dyn_func``: dx/dt of the phase model at one point, states ordered Cn, F (, A, Tau1, Km), the
        controls of the model (last pulse width | the window's pulse intensities) and the window's pulse times as
        ``numerical_timeseries``.  Here it evaluates ``system_dynamics`` on numbers (no casadi graph) and returns a
        ``DynamicsEvaluation(dxdt, defects)`` record like bioptim's; ``parameters`` / ``algebraic_states`` are unused, as in the
        reference."""
        model = fes_model if fes_model is not None else nlp.model
        x = [float(v) for v in np.asarray(states, dtype=float).reshape(-1)]
        u = [] if controls is None else [float(v) for v in np.asarray(controls, dtype=float).reshape(-1)]
        kw = dict(cn=x[0], f=x[1], t=float(np.asarray(time, dtype=float).reshape(-1)[0]),
                  t_stim_prev=[float(v) for v in np.asarray(numerical_timeseries, dtype=float).reshape(-1)],
                  force_length_relationship=force_length_relationship, force_velocity_relationship=force_velocity_relationship,
                  passive_force_relationship=passive_force_relationship)
        if model._with_fatigue:
            kw.update(a=x[2], tau1=x[3], km=x[4])
        if model.kind == FES_DING2007:
            kw["pulse_width"] = u[0]
        elif model.kind == FES_HMED2018:
            kw["pulse_intensity"] = u
        return DynamicsEvaluation(dxdt=model.system_dynamics(**kw), defects=None)

    def declare_ding_variables(self, ocp, nlp, numerical_data_timeseries=None):
        """bioptim's state / control declaration (``StateConfigure`` + ``ConfigureProblem.configure_dynamics_function``) builds
        the casadi NLP, which is the control plane around this path and stays with bioptim: where bioptim
        is installed the reference's own model classes declare the variables and ``cocofest_b200.GpuRK4`` integrates them."""
        raise NotImplementedError(
            "declare_ding_variables assembles bioptim's casadi NLP, which is outside the accelerated path: configure the problem "
            "with the reference's model classes and pass ode_solver=cocofest_b200.GpuRK4(...) (INTEGRATION.md section 3), or "
            "evaluate defects / Jacobians / Hessian blocks directly with cocofest_b200.FesShooting")


class DynamicsEvaluation:
    """The record bioptim's dynamics functions return (``bioptim.DynamicsEvaluation``): ``dxdt`` and ``defects``."""

    def __init__(self, dxdt=None, defects=None):
        self.dxdt = dxdt
        self.defects = defects


class _FatigueMixin:
    """Fatigue states A, Tau1, Km (reference: ding2003_with_fatigue.py:50-59, 197-240)."""

    def _init_fatigue(self):
        d = _fes.default_params(self.kind)
        self.alpha_a = float(d[PARAM_INDEX["alpha_a"]])
        self.alpha_tau1 = float(d[PARAM_INDEX["alpha_tau1"]])
        self.tau_fat = float(d[PARAM_INDEX["tau_fat"]])
        self.alpha_km = float(d[PARAM_INDEX["alpha_km"]])

    def set_alpha_a(self, model, alpha_a):
        self.alpha_a = alpha_a

    def set_alpha_km(self, model, alpha_km):
        self.alpha_km = alpha_km

    def set_alpha_tau1(self, model, alpha_tau1):
        self.alpha_tau1 = alpha_tau1

    def set_tau_fat(self, model, tau_fat):
        self.tau_fat = tau_fat

    def a_dot_fun(self, a, f):
        return -(a - self._a_for_rest()) / self.tau_fat + self.alpha_a * f

    def tau1_dot_fun(self, tau1, f):
        return -(tau1 - self.tau1_rest) / self.tau_fat + self.alpha_tau1 * f

    def km_dot_fun(self, km, f):
        return -(km - self.km_rest) / self.tau_fat + self.alpha_km * f


class DingModelFrequencyWithFatigue(_FatigueMixin, DingModelFrequency):
    """reference: models/ding2003_with_fatigue.py:17"""

    _fatigue = True
    _default_name = "ding2003_with_fatigue"

    def __init__(self, model_name: str = None, muscle_name: str = None, stim_time: list = None,
                 previous_stim: dict = None, sum_stim_truncation: int = 20, is_approximated: bool = False):
        super().__init__(model_name, muscle_name, stim_time, previous_stim, sum_stim_truncation)
        self.is_approximated = is_approximated
        self._init_fatigue()


class DingModelPulseWidthFrequency(DingModelFrequency):
    """Ding 2007: pulse width scales A (reference: models/ding2007.py:17).  As in the reference, the
    constructor's physiological keyword arguments are accepted and ignored (ding2007.py:37-48)."""

    kind = FES_DING2007
    _default_name = "ding_2007"
    _ID_KEYS = ("a_scale", "tau1_rest", "km_rest", "tau2", "pd0", "pdt")

    def __init__(self, model_name: str = None, muscle_name: str = None, stim_time: list = None,
                 previous_stim: dict = None, sum_stim_truncation: int = None, tauc=None, a_rest=None, tau1_rest=None,
                 km_rest=None, tau2=None, pd0=None, pdt=None, a_scale=None, alpha_a=None, alpha_tau1=None,
                 alpha_km=None, tau_fat=None):
        super().__init__(model_name, muscle_name, stim_time, previous_stim, sum_stim_truncation)
        self.pulse_width = None
        d = _fes.default_params(FES_DING2007)
        self.a_scale = float(d[PARAM_INDEX["a_scale"]])
        self.pd0 = float(d[PARAM_INDEX["pd0"]])
        self.pdt = float(d[PARAM_INDEX["pdt"]])
        self.tau1_rest = float(d[PARAM_INDEX["tau1_rest"]])
        self.tau2 = float(d[PARAM_INDEX["tau2"]])
        self.km_rest = float(d[PARAM_INDEX["km_rest"]])
        self.tauc = _fes.DEFAULT_TAUC[FES_DING2007]

    def set_a_scale(self, model, a_scale):
        self.a_scale = a_scale

    def set_pd0(self, model, pd0):
        self.pd0 = pd0

    def set_pdt(self, model, pdt):
        self.pdt = pdt

    def _a_for_rest(self):
        return self.a_scale

    def _serial_keys(self):
        return super()._serial_keys() + ("a_scale", "pd0", "pdt", "stim_time", "previous_stim")

    def a_calculation(self, a_scale, pulse_width):
        return a_scale * (1 - math.exp(-(pulse_width - self.pd0) / self.pdt))

    @staticmethod
    def a_calculation_identification(a_scale, pulse_width, pd0, pdt):
        """``a_calculation`` with pd0 / pdt passed in (the identification OCP's parameters, ding2007.py:190-212; used by
        ``custom_constraints.py:93``)."""
        return a_scale * (1 - math.exp(-(pulse_width - pd0) / pdt))

    def set_impulse_width(self, value):
        """Parameter callback of the pulse-width OCPs (ding2007.py:215-225): the optimised widths become the model's."""
        self.pulse_width = value

    def _a_input(self, a, pulse_width=None, **kw):
        return self.a_calculation(a, pulse_width)


class DingModelPulseWidthFrequencyWithFatigue(_FatigueMixin, DingModelPulseWidthFrequency):
    """reference: models/ding2007_with_fatigue.py:16"""

    _fatigue = True
    _default_name = "ding_2007_with_fatigue"

    def __init__(self, model_name: str = None, muscle_name: str = None, stim_time: list = None,
                 previous_stim: dict = None, sum_stim_truncation: int = 20, **ignored_physiology):
        super().__init__(model_name, muscle_name, stim_time, previous_stim, sum_stim_truncation)
        self._init_fatigue()


class DingModelPulseIntensityFrequency(DingModelFrequency):
    """Hmed 2018: every pulse weighted by lambda_i(I_i) (reference: models/hmed2018.py:17)."""

    kind = FES_HMED2018
    _default_name = "hmed2018"
    _ID_KEYS = ("a_rest", "tau1_rest", "km_rest", "tau2", "ar", "bs", "Is", "cr")

    def __init__(self, model_name: str = None, muscle_name: str = None, stim_time: list = None,
                 previous_stim: dict = None, sum_stim_truncation: int = 20):
        if previous_stim and len(previous_stim["time"]) != len(previous_stim["pulse_intensity"]):
            raise ValueError("The previous_stim time and pulse_intensity must have the same length")
        super().__init__(model_name, muscle_name, stim_time, previous_stim, sum_stim_truncation)
        self.stim_pulse_intensity_prev = []
        self.previous_stim = previous_stim if previous_stim else {"time": [], "pulse_intensity": []}
        d = _fes.default_params(FES_HMED2018)
        self.ar, self.bs = float(d[PARAM_INDEX["ar"]]), float(d[PARAM_INDEX["bs"]])
        self.Is, self.cr = float(d[PARAM_INDEX["Is"]]), float(d[PARAM_INDEX["cr"]])
        self.impulse_intensity = None

    @property
    def pulse_intensity_name(self):
        return "pulse_intensity" + ("_" + self.muscle_name if self.muscle_name else "")

    def set_ar(self, model, ar):
        self.ar = ar

    def set_bs(self, model, bs):
        self.bs = bs

    def set_Is(self, model, Is):
        self.Is = Is

    def set_cr(self, model, cr):
        self.cr = cr

    def _serial_keys(self):
        return super()._serial_keys() + ("ar", "bs", "Is", "cr")

    def lambda_i_calculation(self, pulse_intensity):
        return self.ar * (math.tanh(self.bs * (pulse_intensity - self.Is)) + self.cr)

    @staticmethod
    def lambda_i_calculation_identification(pulse_intensity, ar, bs, Is, cr):
        """``lambda_i_calculation`` with the recruitment parameters passed in (hmed2018.py:183-204; ``custom_constraints.py:41``)."""
        return ar * (math.tanh(bs * (pulse_intensity - Is)) + cr)

    def set_impulse_intensity(self, value):
        """Parameter callback of the pulse-intensity OCPs (hmed2018.py:206-218): one entry per pulse."""
        self.impulse_intensity = [value[i] for i in range(np.shape(value)[0])]

    def get_lambda_i(self, nb_stim: int, pulse_intensity) -> list:
        return [self.lambda_i_calculation(pulse_intensity[i]) for i in range(nb_stim)]

    def min_pulse_intensity(self):
        """Intensity below which lambda_i <= 0 (hmed2018.py:303-310)."""
        return (np.arctanh(-self.cr) / self.bs) + self.Is

    def _get_additional_previous_stim_time(self):
        while len(self.previous_stim["time"]) < self._sum_stim_truncation:
            self.previous_stim["time"].insert(0, _fes.PAD_TIME)
            self.previous_stim["pulse_intensity"].insert(0, _fes.PAD_INTENSITY)
        return self.previous_stim


class DingModelPulseIntensityFrequencyWithFatigue(_FatigueMixin, DingModelPulseIntensityFrequency):
    """reference: models/hmed2018_with_fatigue.py:16"""

    _fatigue = True
    _default_name = "hmed2018_with_fatigue"

    def __init__(self, model_name: str = None, muscle_name: str = None, stim_time: list = None,
                 previous_stim: dict = None, sum_stim_truncation: int = 20):
        super().__init__(model_name, muscle_name, stim_time, previous_stim, sum_stim_truncation)
        self._init_fatigue()


class ModelMaker:
    """String factory (reference: models/model_maker.py:9-22)."""

    _MODELS = {
        "ding2003": DingModelFrequency,
        "ding2003_with_fatigue": DingModelFrequencyWithFatigue,
        "ding2007": DingModelPulseWidthFrequency,
        "ding2007_with_fatigue": DingModelPulseWidthFrequencyWithFatigue,
        "hmed2018": DingModelPulseIntensityFrequency,
        "hmed2018_with_fatigue": DingModelPulseIntensityFrequencyWithFatigue,
    }

    @staticmethod
    def create_model(model_type, **kwargs):
        if model_type not in ModelMaker._MODELS:
            raise ValueError(f"Unknown model type: {model_type}")
        return ModelMaker._MODELS[model_type](**kwargs)
