"""The two step propagators of fes_core.cuh (PROP: scalar, two-state models; PROP2: 2 x 2, reduced fatigue columns) are
algebraic restatements of what RK4 does to the linear tangent rows.  This test compiles the SAME kernel source for the host
twice -- as shipped, and with both propagators switched off (every column through its own four stages,
-DFES_NO_PROPAGATOR -DFES_NO_PROP2) -- and requires the two to agree to rounding on every model and direction set.
(Parity with the oracle is tested elsewhere; this one pins the propagator algebra itself, including the forcing weights.)
"""
import ctypes
import subprocess
from pathlib import Path

import numpy as np
import pytest
import torch

from cocofest_b200 import _lib, fes
from tests import problems as pr

EMUL = Path(__file__).resolve().parent / "host_emul"


def _load(name):
    subprocess.run(["make", "-C", str(EMUL), name], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return _lib._bind(ctypes.CDLL(str(EMUL / name)))


def _run(pb, dtype):
    dev = torch.device("cpu")
    params, x0, win, ctrl = pr.to_lib(pb, dev, dtype)
    out = {}
    for dirs, tag in ((fes.FES_DIR_X0, "x0"), (fes.FES_DIR_P, "p")):
        r = fes.ivp(pb["kind"], pb["fat"], params, x0, win, pb["dt"], 10, ctrl=ctrl, dirs=dirs)
        out["ivp_" + tag] = [r.x_final, r.sens_final]
    rng = np.random.default_rng(11)
    X = torch.tensor(rng.uniform(0.9, 1.1, (pb["ns"] + 1, pb["nx"], pb["B"])) * pb["x0"].T[None], dtype=dtype)
    wants = [("x",), ("w",), ("p",)] + ([("x", "u"), ("u",)] if pb["nu"] else [])
    for want in wants:
        s = fes.shoot(pb["kind"], pb["fat"], params, X, win, pb["dt"], 10, ctrl=ctrl, want=want)
        out["shoot_" + "".join(want)] = [t for t in (s.G, s.Jx, s.Ju, s.Jp, s.Jw) if t is not None]
    return {k: [t.double().numpy().copy() for t in v] for k, v in out.items()}


@pytest.mark.parametrize("name", list(pr.MODELS))
@pytest.mark.parametrize("dtype,tol", [(torch.float64, 2e-13), (torch.float32, 2e-4)])
def test_propagators_equal_stage_by_stage(name, dtype, tol, monkeypatch):
    pb = pr.random_problem(name, 6, N=6, ns=7, seed=3)
    res = {}
    for lib in ("libfes_emul.so", "libfes_emul_stages.so"):
        monkeypatch.setattr(_lib, "_LIB", _load(lib))
        res[lib] = _run(pb, dtype)
    differs = False
    for key, arrs in res["libfes_emul.so"].items():
        for a, b in zip(arrs, res["libfes_emul_stages.so"][key]):
            assert a.shape == b.shape
            assert pr.rel_err(a, b, floor=1e-9 if dtype == torch.float64 else 1e-4) < tol, key
            differs |= not np.array_equal(a, b)
    assert differs  # the two builds really took different code paths
