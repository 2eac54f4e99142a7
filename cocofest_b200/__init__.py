"""cocofest_b200 -- B200-native batched integrator for cocofest's FES muscle models.

Host side (Python/PyTorch) mirrors the reference surface for the hot path only:
``DingModelFrequency[WithFatigue]``, ``DingModelPulseWidthFrequency[WithFatigue]``,
``DingModelPulseIntensityFrequency[WithFatigue]``, ``ModelMaker``, ``IvpFes``, ``OcpFes.prepare_n_shooting``
and the multiple-shooting defect / Jacobian evaluation; the arithmetic runs in hand-written sm_100a CUDA
kernels behind the C ABI of ``include/fes_b200.h``.
"""
from . import fes, hill_coefficients
from .models import (
    DingModelFrequency,
    DingModelFrequencyWithFatigue,
    DingModelPulseIntensityFrequency,
    DingModelPulseIntensityFrequencyWithFatigue,
    DingModelPulseWidthFrequency,
    DingModelPulseWidthFrequencyWithFatigue,
    FesModel,
    ModelMaker,
)
from .ode_solver import OdeSolver
from .ivp_fes import IvpFes
from .ocp_fes import OcpFes, FesShooting, FesNmpcBatch
from .bioptim_seam import FesIntegrator, GpuRK4
from .custom_objectives import CustomObjective
from .pickle_io import SolutionToPickle
from .fourier_approx import FourierSeries
from .identification import (
    DingModelFrequencyForceParameterIdentification,
    DingModelPulseIntensityFrequencyForceParameterIdentification,
    DingModelPulseWidthFrequencyForceParameterIdentification,
)

__all__ = [
    "fes", "hill_coefficients", "FesModel", "ModelMaker", "OdeSolver", "IvpFes", "OcpFes", "FesShooting", "FesNmpcBatch", "FesIntegrator", "GpuRK4", "FourierSeries", "CustomObjective", "SolutionToPickle",
    "DingModelFrequency", "DingModelFrequencyWithFatigue",
    "DingModelPulseWidthFrequency", "DingModelPulseWidthFrequencyWithFatigue",
    "DingModelPulseIntensityFrequency", "DingModelPulseIntensityFrequencyWithFatigue",
    "DingModelFrequencyForceParameterIdentification", "DingModelPulseWidthFrequencyForceParameterIdentification",
    "DingModelPulseIntensityFrequencyForceParameterIdentification",
]
