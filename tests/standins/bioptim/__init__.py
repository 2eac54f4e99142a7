"""A STAND-IN for the one bioptim name cocofest_b200.bioptim_seam.GpuRK4 touches (`OdeSolver.RK4`, ode_solver.py:186-271 of
bioptim 3.2.1) -- TEST INFRASTRUCTURE ONLY: a solver object that keeps `n_integration_steps` and exposes the integrator CLASS
through the `integrator` property, which bioptim instantiates as `nlp.ode_solver.integrator(ode, ode_opt)` (ode_solver.py:232-271)."""


class _RK:
    def __init__(self, n_integration_steps: int = 5, **kwargs):
        self.n_integration_steps = n_integration_steps

    @property
    def integrator(self):
        raise NotImplementedError("stand-in: the stock RK integrator is casadi code")

    def initialize_integrator(self, ode, ode_opt):
        """What bioptim does with the solver for one phase (ode_solver.py:232-271, condensed)."""
        return self.integrator(ode, ode_opt)


class OdeSolver:
    class RK1(_RK):
        pass

    class RK2(_RK):
        pass

    class RK4(_RK):
        pass
