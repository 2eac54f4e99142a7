"""The reference's ``CustomObjective`` (cocofest/custom_objectives.py) for B scenarios at once.

In the reference these are bioptim penalty closures over a ``PenaltyController`` (one node of one casadi problem).  Here the
same quantities are evaluated on the node-state tensors ``FesShooting`` works with -- ``X [ns+1, nx, B]`` on the device --
and every method returns ``(cost [B], d cost / d X)`` with the weights, node sets and quadratic form that
``OcpFes._set_objective`` / ``OcpFesMsk`` give them (fes_ocp.py:530-571, fes_ocp_dynamics.py:620-703): what a solver needs
beside the defects and Jacobian blocks.  Same method names; the first argument is the tensor instead of the controller."""
from __future__ import annotations

import numpy as np

from .fourier_approx import (FourierSeries, best_fit_objective, muscle_fatigue_objective, muscle_force_objective,
                             tracking_objective)


class CustomObjective:
    @staticmethod
    def track_state_from_time(X, dt, fourier_coeff, key="F", weight=100.0):
        """custom_objectives.py:15-40: the Fourier series of the target evaluated at the node times (period 1, nodes at
        ``linspace(0, 1, ns+1)`` as fes_ocp.py:539-548 samples it) minus the state, squared, Lagrange over all nodes."""
        row = {"Cn": 0, "F": 1, "A": 2, "Tau1": 3, "Km": 4}[key]
        target = FourierSeries().fit_func_by_fourier_series_with_real_coeffs(np.linspace(0, 1, X.shape[0]), fourier_coeff)
        return tracking_objective(X, dt, target=target, weight=weight, force_row=row)

    @staticmethod
    def track_state_from_time_interpolate(X, dt, force, key="F", minimization_type="least square", weight=1.0):
        """custom_objectives.py:43-74: ``force`` [ns+1(, B)] at the nodes against the state, least square or best fit."""
        row = {"Cn": 0, "F": 1, "A": 2, "Tau1": 3, "Km": 4}[key]
        if minimization_type == "least square":
            return tracking_objective(X, dt, target=force, weight=weight, force_row=row)
        if minimization_type == "best fit":
            return best_fit_objective(X, dt, force, weight=weight, force_row=row)
        raise RuntimeError(f"Minimization type {minimization_type} not implemented")

    @staticmethod
    def minimize_overall_muscle_fatigue(X_end, a_rest, weight=1.0):
        """custom_objectives.py:77-97 (Mayer, last node): (sum of a_rest / sum of A over the muscles)^2."""
        return muscle_fatigue_objective(X_end, a_rest, weight=weight)

    @staticmethod
    def minimize_overall_muscle_force_production(X, dt, weight=1.0):
        """custom_objectives.py:100-117 (Lagrange, all nodes): sum of the squared muscle forces."""
        return muscle_force_objective(X, dt, weight=weight)
