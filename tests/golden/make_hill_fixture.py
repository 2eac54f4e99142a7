"""Golden fixture for the Hill-type relationships: executes the REFERENCE'S OWN functions
(cocofest/models/hill_coefficients.py) in the build container and writes hill_golden.json.

    python tests/golden/make_hill_fixture.py      (needs /root/reference)

The reference module imports exp / log / sqrt from casadi and asks a biorbd muscle for its length, velocity and optimal
length; neither library is installed, so ``casadi`` is a stand-in whose three functions are numpy's and the muscle is a
stand-in that returns given numbers.  The function bodies run unchanged."""
import importlib.util
import json
import sys
import types
from pathlib import Path

import numpy as np

REF = "/root/reference/cocofest/models/hill_coefficients.py"
OUT = Path(__file__).with_name("hill_golden.json")


class _Num:
    def __init__(self, v):
        self.v = v

    def to_mx(self):
        return self.v


class _Muscle:
    """Stands in for a biorbd muscle: length(model, q, update), velocity(model, q, qdot, update), characteristics()."""

    def __init__(self, length, velocity, optimal):
        self._l, self._v, self._o = length, velocity, optimal

    def length(self, model, q, update):
        return _Num(self._l)

    def velocity(self, model, q, qdot, update):
        return _Num(self._v)

    def characteristics(self):
        return self

    def optimalLength(self):
        return _Num(self._o)


def main():
    cas = types.ModuleType("casadi")
    cas.exp, cas.log, cas.sqrt = np.exp, np.log, np.sqrt
    sys.modules["casadi"] = cas
    spec = importlib.util.spec_from_file_location("ref_hill", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    optimal = 0.13
    rows = []
    for nl in np.linspace(0.45, 1.65, 25):
        for v in (-6.0, -1.5, -0.2, 0.0, 0.3, 2.0, 7.5):
            m = _Muscle(float(nl * optimal), float(v), optimal)
            rows.append({"norm_length": float(nl), "velocity": float(v),
                         "fl": float(mod.muscle_force_length_coefficient(None, m, None)),
                         "fv": float(mod.muscle_force_velocity_coefficient(None, m, None, None)),
                         "fp": float(mod.muscle_passive_force_coefficient(None, m, None))})
    OUT.write_text(json.dumps({"how": "reference hill_coefficients.py executed from /root/reference with stand-ins for casadi and the biorbd muscle",
                               "optimal_length": optimal, "rows": rows}))
    print("wrote", OUT, len(rows), "rows")


if __name__ == "__main__":
    main()
