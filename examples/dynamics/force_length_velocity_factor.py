"""The force-length / force-velocity / passive-force coupling of the FES-MSK models on the batched path.

Reference: examples/dynamics/muscle_force_length_and_force_velocity_relationships_comparison.py compares the joint angle of
an arm model with and without the relationships; there the three coefficients are evaluated from the joint states inside the
muscle ODE (f_dot_fun, ding2003.py:309-311; hill_coefficients.py).  The rigid-body model (biorbd) is not part of this package:
here the joint motion is PRESCRIBED -- every scenario flexes the elbow along its own smooth trajectory, given as normalised
fibre length and fibre velocity at the nodes -- and the muscle side is what runs on the GPU:

  1. ``hill_coefficients.force_factor`` turns (length, velocity) of every (node, scenario) into fl * fv + fp and its
     derivatives;
  2. ``FesShooting.evaluate(..., force_factor=)`` integrates every interval with its factor and returns the end points and
     dPhi/dx, dPhi/du, dPhi/dfactor;
  3. the chain rule gives dPhi/d(length), dPhi/d(velocity) -- what a solver that owns the joint states needs -- checked here
     against differences of the forward map.

    python examples/dynamics/force_length_velocity_factor.py [--scenarios 20000]
"""
import argparse
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))

from cocofest_b200 import DingModelPulseWidthFrequencyWithFatigue, FesShooting, OdeSolver, hill_coefficients  # noqa: E402


def forward_forces(shooting, x0, U, factor):
    """Single shooting on the interval kernel: node states chained through the end points (factor [ns, B] or None)."""
    ns, B = shooting.n_shooting, x0.shape[1]
    X = x0[None].repeat(ns + 1, 1, 1).contiguous()
    for k in range(ns):   # interval k only needs x_k: evaluate all, keep the k-th end point (cheap at this size)
        res = shooting.evaluate(X, U, want=("x",), defects=False, endpoints=True, force_factor=factor)
        X[k + 1] = res.Phi[k]
    return X


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scenarios", type=int, default=20000)
    ap.add_argument("--device", default="cuda")
    args = ap.parse_args()
    dev, B = torch.device(args.device), args.scenarios
    model = DingModelPulseWidthFrequencyWithFatigue(stim_time=list(np.linspace(0, 1, 11)[:-1]), sum_stim_truncation=10)
    shooting = FesShooting(model, final_time=1, ode_solver=OdeSolver.RK4(n_integration_steps=10), device=dev)
    ns = shooting.n_shooting
    gen = torch.Generator().manual_seed(0)
    # prescribed joint motion per scenario: the fibre shortens from l0 to l1 along a raised cosine over the second
    l0 = (1.05 + 0.25 * torch.rand(B, generator=gen, dtype=torch.float64)).to(dev)
    l1 = (0.75 + 0.20 * torch.rand(B, generator=gen, dtype=torch.float64)).to(dev)
    t_mid = ((torch.arange(ns, dtype=torch.float64) + 0.5) / ns).to(dev)[:, None]          # the factor is held at the interval's midpoint
    s = 0.5 * (1 - torch.cos(torch.pi * t_mid))
    length = l0 + (l1 - l0) * s                                                             # [ns, B] normalised fibre length
    velocity = (l1 - l0) * 0.5 * torch.pi * torch.sin(torch.pi * t_mid) * 0.13              # [ns, B] m/s for a 13 cm optimal length
    factor, dfac_dl, dfac_dv = hill_coefficients.force_factor(length, velocity, derivative=True)
    x0 = torch.tensor(model.standard_rest_values(), dtype=torch.float64, device=dev).reshape(-1, 1).repeat(1, B).contiguous()
    U = torch.full((ns, 1), 0.00025, dtype=torch.float64, device=dev)                       # pulse_width = {"fixed": 0.00025}

    t0 = time.perf_counter()
    X_with = forward_forces(shooting, x0, U, factor)
    X_without = forward_forces(shooting, x0, U, None)
    torch.cuda.synchronize() if dev.type == "cuda" else None
    seconds = time.perf_counter() - t0
    f_with, f_without = X_with[-1, 1], X_without[-1, 1]
    print(f"{B} scenarios x {ns} intervals, with and without the relationships: {seconds:.3f} s")
    print(f"  force at t = 1 s without relationships {f_without.mean().item():8.2f} N (the same for every scenario: {f_without.std().item():.1e})")
    print(f"  force at t = 1 s with    relationships {f_with.mean().item():8.2f} +- {f_with.std().item():.2f} N, factor in "
          f"[{factor.min().item():.3f}, {factor.max().item():.3f}]")

    # the Jacobian blocks of the multiple-shooting problem at these node states, and the chain rule to the joint side
    res = shooting.evaluate(X_with, U, endpoints=True, force_factor=factor)
    dPhi_dl = res.Jc * dfac_dl[:, None, :]       # [ns, nx, B]
    dPhi_dv = res.Jc * dfac_dv[:, None, :]
    h = 1e-6
    up = shooting.evaluate(X_with, U, want=("x",), defects=False, endpoints=True,
                           force_factor=hill_coefficients.force_factor(length + h, velocity)).Phi
    dn = shooting.evaluate(X_with, U, want=("x",), defects=False, endpoints=True,
                           force_factor=hill_coefficients.force_factor(length - h, velocity)).Phi
    fd = (up - dn) / (2 * h)
    err = ((dPhi_dl - fd).abs().amax(dim=(0, 2)) / fd.abs().amax(dim=(0, 2)).clamp_min(1e-300)).max().item()
    print(f"  d Phi / d fibre length through Jc: max relative deviation from central differences {err:.1e}")
    assert err < 1e-4 and float(res.G.abs().max()) < 1e-9 * float(X_with.abs().max()) + 1e-9
    assert float((f_with - f_without).abs().max()) > 1.0, "the relationships must change the force"
    return {"force_with": f_with, "force_without": f_without, "dPhi_dlength": dPhi_dl, "dPhi_dvelocity": dPhi_dv, "fd_error": err}


if __name__ == "__main__":
    main()
