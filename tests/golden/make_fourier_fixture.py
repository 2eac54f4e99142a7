"""Golden fixture for the force-tracking target: executes the REFERENCE'S OWN ``FourierSeries``
(cocofest/fourier_approx.py) in the build container and writes fourier_golden.json.

    python tests/golden/make_fourier_fixture.py      (needs /root/reference and scipy)

The reference module imports casadi (cos, sin -- only used in its "casadi" mode) and matplotlib at module top;
neither is installed, so inert stand-ins are put in ``sys.modules`` and the numpy mode is run unchanged."""
import importlib.util
import json
import sys
import types
from pathlib import Path

import numpy as np

REF = "/root/reference/cocofest/fourier_approx.py"
OUT = Path(__file__).with_name("fourier_golden.json")


def main():
    cas = types.ModuleType("casadi")
    cas.cos, cas.sin = np.cos, np.sin
    sys.modules["casadi"] = cas
    mpl = types.ModuleType("matplotlib")
    plt = types.ModuleType("matplotlib.pyplot")
    mpl.pyplot = plt
    sys.modules["matplotlib"], sys.modules["matplotlib.pyplot"] = mpl, plt
    spec = importlib.util.spec_from_file_location("ref_fourier", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    fs = mod.FourierSeries()
    # the force profile of examples/getting_started/force_tracking_parameter_optimization.py: a sampled, smooth,
    # non-periodic-looking curve on [0, 1]
    x = np.linspace(0, 1, 101)
    y = 30 + 150 * np.abs(np.sin(np.pi * x)) ** 1.5 + 10 * x
    cases = []
    for n, ns in ((50, 10), (50, 33), (8, 100)):
        ab = fs.compute_real_fourier_coeffs(x, y, n)
        nodes = np.linspace(0, 1, ns + 1)
        val = fs.fit_func_by_fourier_series_with_real_coeffs(nodes, ab)
        cases.append({"n": n, "n_shooting": ns, "ab": ab.tolist(), "force_to_track": np.asarray(val).tolist()})
    OUT.write_text(json.dumps({"how": "reference FourierSeries (numpy mode) executed from /root/reference",
                               "x": x.tolist(), "y": y.tolist(), "cases": cases}))
    print("wrote", OUT)


if __name__ == "__main__":
    main()
