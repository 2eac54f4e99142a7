"""Minimal stand-in for bioptim's ``OdeSolver`` namespace (external/bioptim/bioptim/dynamics/ode_solver.py):
only what the FES path reads -- the scheme and ``n_integration_steps`` -- so that
``OdeSolver.RK4(n_integration_steps=10)`` keeps working in user scripts.  Objects exposing the same two
attributes (e.g. a real bioptim OdeSolver where bioptim is installed) are accepted as well."""


class _RK:
    scheme = None

    def __init__(self, n_integration_steps: int = 5):
        if not isinstance(n_integration_steps, int) or n_integration_steps < 1:
            raise ValueError("n_integration_steps must be a positive int")
        self.n_integration_steps = n_integration_steps

    def __repr__(self):
        return f"OdeSolver.{type(self).__name__}(n_integration_steps={self.n_integration_steps})"


class OdeSolver:
    class RK1(_RK):
        scheme = 1

    class RK2(_RK):
        scheme = 2

    class RK4(_RK):
        scheme = 4

    class COLLOCATION:
        """Accepted by the reference's type check (ivp_fes.py:223-227) but not part of the GPU path."""

        def __init__(self, *a, **k):
            raise NotImplementedError("collocation is outside the accelerated path; use RK1/RK2/RK4")


def scheme_and_steps(ode_solver):
    """(scheme, n_steps) of an OdeSolver-like object."""
    scheme = getattr(ode_solver, "scheme", None)
    if scheme is None:
        scheme = {"RK1": 1, "RK2": 2, "RK4": 4}.get(type(ode_solver).__name__)
    steps = getattr(ode_solver, "n_integration_steps", None)
    if steps is None:
        steps = getattr(ode_solver, "steps", None)
    if scheme not in (1, 2, 4) or not isinstance(steps, int):
        raise ValueError("ode_solver must be a OdeSolver type")
    return scheme, steps
