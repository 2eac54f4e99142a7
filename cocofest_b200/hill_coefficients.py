"""Hill-type muscle relationships of De Groote et al. (2016), as the reference uses them for its FES-MSK models
(cocofest/models/hill_coefficients.py:11-126), written for arrays of normalised fibre lengths / fibre velocities instead of a
biorbd muscle and casadi symbols, with their derivatives.

They produce the force factor ``fl * fv + fp`` of ``f_dot_fun`` (ding2003.py:309-311) that ``fes.shoot(..., force_factor=)``
takes per (scenario, interval); the derivatives chain ``ShootResult.Jc`` to the joint states of the caller's rigid-body model:

    d Phi / d q = Jc * (d factor / d norm_length * d norm_length / d q + d factor / d velocity * d velocity / d q)

The rigid-body side (muscle lengths and velocities from q, qdot: biorbd in the reference) is out of scope here.
Works on numpy arrays and torch tensors alike (only +, *, /, exp, log, sqrt and a clamp are used).
"""
import math

# (amplitude, centre, width offset, width slope) of the three gaussians of the active force-length curve
_FL_TERMS = ((0.815, 1.055, 0.162, 0.063), (0.433, 0.717, -0.030, 0.200), (0.100, 1.000, 0.354, 0.0))
_FV = (-0.318, -8.149, -0.374, 0.886)   # d1 .. d4 of the force-velocity curve
MAX_SHORTENING_SPEED = 10.0             # the reference normalises the fibre velocity by this constant (hill_coefficients.py:85)
_KPE, _E0 = 4.0, 0.6                    # passive curve


def _ns(x):
    """The array namespace of x (torch or numpy) for exp / log / sqrt."""
    try:
        import torch

        if isinstance(x, torch.Tensor):
            return torch
    except ImportError:  # pragma: no cover
        pass
    import numpy

    return numpy


def muscle_force_length_coefficient(norm_length, derivative=False):
    """Active force-length factor at the normalised fibre length l = length / optimal length (hill_coefficients.py:11-61).
    derivative=True: also d / d l."""
    xp = _ns(norm_length)
    l = norm_length
    val, der = 0.0, 0.0
    for amp, centre, w0, w1 in _FL_TERMS:
        num, den = l - centre, w0 + w1 * l
        g = amp * xp.exp(-0.5 * (num * num) / (den * den))
        val = val + g
        # d/dl of -(num^2)/(2 den^2) = -num/den^2 + num^2 w1/den^3
        der = der + g * (-(num / (den * den)) + (num * num) * w1 / (den * den * den))
    return (val, der) if derivative else val


def muscle_force_velocity_coefficient(muscle_velocity, derivative=False):
    """Force-velocity factor at the fibre velocity v (same unit as the reference's ``muscle.velocity``); the curve is taken at
    v / MAX_SHORTENING_SPEED (hill_coefficients.py:64-95).  derivative=True: also d / d v."""
    xp = _ns(muscle_velocity)
    d1, d2, d3, d4 = _FV
    z = d2 * (muscle_velocity / MAX_SHORTENING_SPEED) + d3
    root = xp.sqrt(z * z + 1.0)
    val = d1 * xp.log(z + root) + d4           # d1 asinh(z) + d4
    if not derivative:
        return val
    return val, d1 * (d2 / MAX_SHORTENING_SPEED) / root


def muscle_passive_force_coefficient(norm_length, derivative=False):
    """Passive factor at the normalised fibre length, zero below the optimal length (hill_coefficients.py:98-126).
    derivative=True: also d / d l (zero where the factor is clamped)."""
    xp = _ns(norm_length)
    scale = 1.0 / (math.exp(_KPE) - 1.0)
    e = xp.exp((_KPE / _E0) * (norm_length - 1.0))
    raw = (e - 1.0) * scale
    on = raw > 0
    val = raw * on
    if not derivative:
        return val
    return val, (_KPE / _E0) * e * scale * on


def force_factor(norm_length=None, muscle_velocity=None, force_length=True, force_velocity=True, passive_force=True, derivative=False):
    """``fl * fv + fp`` with the relationships that are switched on (the reference's ``activate_*_relationship`` flags,
    dynamical_model.py:36-38; a relationship that is off contributes its neutral value 1 / 1 / 0).
    derivative=True: returns (factor, d factor / d norm_length, d factor / d muscle_velocity)."""
    need_l = force_length or passive_force
    if need_l and norm_length is None:
        raise ValueError("norm_length is needed for the force-length and passive relationships")
    if force_velocity and muscle_velocity is None:
        raise ValueError("muscle_velocity is needed for the force-velocity relationship")
    fl, dfl = muscle_force_length_coefficient(norm_length, True) if force_length else (1.0, 0.0)
    fv, dfv = muscle_force_velocity_coefficient(muscle_velocity, True) if force_velocity else (1.0, 0.0)
    fp, dfp = muscle_passive_force_coefficient(norm_length, True) if passive_force else (0.0, 0.0)
    c = fl * fv + fp
    if not derivative:
        return c
    return c, dfl * fv + dfp, fl * dfv
