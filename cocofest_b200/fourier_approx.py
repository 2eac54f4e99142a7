"""Force-tracking objective of ``OcpFes`` for B scenarios at once.

Reference: ``cocofest/fourier_approx.py:7-38`` (real Fourier coefficients of a sampled target by the trapezoid
rule, period 1; evaluation ``a0/2 + sum a_n cos(2 pi n x) + b_n sin(2 pi n x)``) and ``OcpFes._set_objective``
(``optimization/fes_ocp.py:530-571``): 50 harmonics (``_build_fourier_coefficient``, :347-348), the series
sampled at ``linspace(0, 1, n_shooting + 1)``, tracked by ``Lagrange.TRACK_STATE`` on ``F`` at every node with
weight 100 (quadratic), plus ``Mayer.MINIMIZE_STATE`` on the last node for ``end_node_tracking`` (weight 1).

The coefficient fit is host numpy (a 51 x len(target) product, once per problem).  The cost and its gradient with
respect to the node forces are elementwise operations on the ``[ns+1, nx, B]`` device tensor that ``FesShooting``
evaluates; the node weight is ``dt`` as in the identification objective (fes_ident.cuh), so the two agree.
"""
from __future__ import annotations

import numpy as np
import torch


class FourierSeries:
    def __init__(self):
        self.p = 1  # period value

    def compute_real_fourier_coeffs(self, x, y, n):
        """(a_i, b_i), i = 0..n, of the samples (x, y) by the trapezoid rule (fourier_approx.py:12-18)."""
        x, y = np.asarray(x, dtype=float), np.asarray(y, dtype=float)
        i = np.arange(n + 1)[:, None]
        w = 2 * np.pi * i * x[None, :] / self.p
        an = (2.0 / self.p) * np.trapezoid(y[None, :] * np.cos(w), x, axis=1)
        bn = (2.0 / self.p) * np.trapezoid(y[None, :] * np.sin(w), x, axis=1)
        return np.stack([an, bn], axis=1)

    def fit_func_by_fourier_series_with_real_coeffs(self, x, ab, mode="numpy"):
        """Value of the series at ``x`` (fourier_approx.py:21-38); ``mode`` is accepted for signature parity."""
        if mode not in ("numpy", "casadi"):
            raise ValueError("mode must be 'numpy' or 'casadi'")
        x = np.asarray(x, dtype=float)
        ab = np.asarray(ab, dtype=float)
        n = np.arange(1, len(ab)).reshape((-1,) + (1,) * x.ndim)
        a = ab[1:, 0].reshape(n.shape)
        b = ab[1:, 1].reshape(n.shape)
        w = 2.0 * np.pi * n * x[None] / self.p
        return ab[0, 0] / 2.0 + np.sum(a * np.cos(w) + b * np.sin(w), axis=0)


    def fourier_approx(self, x, y, n, show=False):
        """The approximation of the samples by their first ``n`` harmonics, evaluated at ``x`` (fourier_approx.py:40-49).  The
        reference also draws both curves with matplotlib; plotting is outside this path, so that part is opt-in (``show=True``)
        and needs matplotlib."""
        ab = self.compute_real_fourier_coeffs(x, y, n)
        y_approx = self.fit_func_by_fourier_series_with_real_coeffs(x, ab)
        if show:
            import matplotlib.pyplot as plt

            plt.scatter(x, y, color="blue", s=5, marker=".")
            plt.plot(x, y_approx, color="red", linewidth=1)
            plt.show()
        return y_approx


def force_to_track(force_tracking, n_shooting, n_harmonics=50):
    """fes_ocp.py:539-548: the target force at the ns+1 nodes from ``force_tracking = [time, force]``."""
    fs = FourierSeries()
    ab = fs.compute_real_fourier_coeffs(force_tracking[0], force_tracking[1], n_harmonics)
    return fs.fit_func_by_fourier_series_with_real_coeffs(np.linspace(0, 1, n_shooting + 1), ab)


def tracking_objective(X, dt, target=None, end_node_target=None, weight=100.0, end_weight=1.0, force_row=1):
    """Cost [B] and its gradient d cost / d X [ns+1, nx, B] of the OcpFes force objectives for B scenarios.

    X [ns+1, nx, B] node states (device tensor); target [ns+1] | [ns+1, B] | None (Lagrange tracking of F, weight
    ``weight``, node weight dt); end_node_target scalar | [B] | None (Mayer term on the last node)."""
    F = X[:, force_row]
    cost = torch.zeros(X.shape[-1], dtype=X.dtype, device=X.device)
    grad = torch.zeros_like(X)
    if target is not None:
        t = torch.as_tensor(target, dtype=X.dtype, device=X.device)
        if t.dim() == 1:
            t = t[:, None]
        if t.shape[0] != F.shape[0]:
            raise ValueError(f"target must have {F.shape[0]} rows (one per node)")
        r = F - t
        cost = cost + (weight * dt) * (r * r).sum(dim=0)
        grad[:, force_row] += (2.0 * weight * dt) * r
    if end_node_target is not None:
        e = F[-1] - torch.as_tensor(end_node_target, dtype=X.dtype, device=X.device)
        cost = cost + end_weight * e * e
        grad[-1, force_row] += 2.0 * end_weight * e
    return cost, grad


def best_fit_objective(X, dt, target, weight=1.0, force_row=1):
    """``track_state_from_time_interpolate(..., minimization_type="best fit")`` (custom_objectives.py:43-74): residual
    ``1 - target / F`` at every node, squared, node weight dt.  Cost [B] and gradient [ns+1, nx, B]."""
    F = X[:, force_row]
    t = torch.as_tensor(target, dtype=X.dtype, device=X.device)
    if t.dim() == 1:
        t = t[:, None]
    if t.shape[0] != F.shape[0]:
        raise ValueError(f"target must have {F.shape[0]} rows (one per node)")
    r = 1.0 - t / F
    grad = torch.zeros_like(X)
    grad[:, force_row] = (2.0 * weight * dt) * r * (t / (F * F))
    return (weight * dt) * (r * r).sum(dim=0), grad


def muscle_fatigue_objective(X_end, a_rest, weight=1.0, a_row=2):
    """``minimize_overall_muscle_fatigue`` as ``fes_ocp_dynamics.py:675-683`` uses it (Mayer, last node, quadratic, weight 1):
    ``(sum_m a_rest_m / sum_m A_m(T))^2`` over the muscles of a scenario (custom_objectives.py:77-97).

    X_end [M, nx, B] end states of the M muscles (or [nx, B] for one), a_rest [M] | [M, B].  Cost [B], gradient like X_end."""
    single = X_end.dim() == 2
    Xe = X_end[None] if single else X_end
    ar = torch.as_tensor(a_rest, dtype=Xe.dtype, device=Xe.device).reshape(Xe.shape[0], -1)
    num = ar.sum(dim=0) if ar.shape[1] > 1 else ar.sum()
    den = Xe[:, a_row].sum(dim=0)
    ratio = num / den
    grad = torch.zeros_like(Xe)
    grad[:, a_row] = (-2.0 * weight * ratio * ratio / den)[None].expand(Xe.shape[0], -1)
    return weight * ratio * ratio, (grad[0] if single else grad)


def muscle_force_objective(X, dt, weight=1.0, force_row=1):
    """``minimize_overall_muscle_force_production`` as ``fes_ocp_dynamics.py:685-693`` uses it (Lagrange, all nodes, quadratic,
    weight 1): ``sum_nodes dt sum_m F_m^2`` (custom_objectives.py:100-117).

    X [M, ns+1, nx, B] node states of the M muscles (or [ns+1, nx, B] for one).  Cost [B], gradient like X."""
    single = X.dim() == 3
    Xm = X[None] if single else X
    F = Xm[:, :, force_row]
    grad = torch.zeros_like(Xm)
    grad[:, :, force_row] = (2.0 * weight * dt) * F
    return (weight * dt) * (F * F).sum(dim=(0, 1)), (grad[0] if single else grad)
