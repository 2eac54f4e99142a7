"""The evaluation loop of a cyclic NMPC for MANY scenarios at once.

Reference: examples/getting_started/pulse_duration_optimization_nmpc_cyclic.py builds ONE NmpcFes problem
(DingModelPulseWidthFrequencyWithFatigue, 33 pulses per 1 s cycle, truncation 10) and lets Ipopt call the casadi graph of the
continuity defects and their Jacobian once per iteration; between cycles FesNmpc.update_stim carries the last pulses over and
the solved horizon is shifted (fes_nmpc.py:44-85).  The solver stays on the host and is not part of this path.  What this
script shows is everything the solver CONSUMES per iteration, for 16 384 scenarios in one call:

    defects g_k = x_{k+1} - Phi_k(x_k, u_k),  dPhi_k/dx_k,  dPhi_k/du_k          fes.ShootPlan (one call per iteration)
    exact Hessian blocks sum_r lambda_r d2Phi_r/dz dz of the continuity rows          FesShooting.hessian (one launch per iteration)
    window carry-over and warm-start shift between cycles                         FesNmpcBatch.update_stim / .advance

and checks the one property that needs no solver: node states produced by integrating a control sequence are defect-free.

    python examples/getting_started/nmpc_cyclic_batch.py [--scenarios 16384] [--cycles 3]
"""
import argparse
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))

from cocofest_b200 import DingModelPulseWidthFrequencyWithFatigue, OdeSolver, fes  # noqa: E402
from cocofest_b200.ocp_fes import FesNmpcBatch  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--scenarios", type=int, default=16384)
ap.add_argument("--cycles", type=int, default=3)
args = ap.parse_args()
B, dev = args.scenarios, torch.device("cuda")

model = DingModelPulseWidthFrequencyWithFatigue(stim_time=[k / 33 for k in range(33)], sum_stim_truncation=10)
nmpc = FesNmpcBatch(model, cycle_duration=1.0, n_cycles_simultaneous=1, n_cycles_to_advance=1,
                    ode_solver=OdeSolver.RK4(n_integration_steps=10), device=dev)
ns = nmpc.n_shooting
params = torch.tensor(model.param_vector(), device=dev)
gen = torch.Generator(device="cpu").manual_seed(0)
pulse_width = (model.pd0 + (6e-4 - model.pd0) * torch.rand((ns, 1, B), generator=gen, dtype=torch.float64)).to(dev)
x = torch.zeros((5, B), dtype=torch.float64, device=dev)
x[2], x[3], x[4] = params[0], params[2], params[1]       # rest: A = a_scale, Tau1 = tau1_rest, Km = km_rest

for cycle in range(args.cycles):
    shooting = nmpc.shooting
    # a feasible point of this horizon: integrate the current controls from the current state (what a warm start is)
    run = fes.ivp(model.kind, True, params, x, shooting.window, shooting.dt, n_steps=10, ctrl=pulse_width, n_shooting=ns, trace="nodes")
    X = run.trace.contiguous()
    plan = shooting.plan(B, pulse_width, want=("x", "u"))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20):                                   # 20 "solver iterations" on the same horizon
        res = plan(X, pulse_width)
    torch.cuda.synchronize()
    per_iter = (time.perf_counter() - t0) / 20
    worst = float((res.G.abs() / X[1:].abs().amax(dim=(0, 2), keepdim=True)).max())
    # what Ipopt's default (exact Hessian, ipopt_options.py:88) adds per iteration: the blocks of the Lagrangian Hessian of the
    # continuity rows for the current multipliers, from the second-order tangent kernel; the full tensor for a few scenarios
    lam = torch.randn((ns, 5, B), generator=gen, dtype=torch.float64).to(dev)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    hes = shooting.hessian(X, pulse_width, multipliers=lam, second=False)
    torch.cuda.synchronize()
    t_hes = time.perf_counter() - t0
    D2 = shooting.second_derivatives(X[:, :, :256].contiguous(), pulse_width[:, :, :256].contiguous())
    Hd = torch.einsum("krb,krijb->kijb", lam[:, :, :256], D2)
    assert float((hes.full(hes.H)[..., :256] - Hd).abs().max()) <= 1e-12 * float(Hd.abs().max())
    print(f"cycle {cycle}: {B} scenarios x {ns} intervals, defects + dPhi/dx + dPhi/dpw in {1e3 * per_iter:.3f} ms per iteration "
          f"({B * ns / per_iter:.3g} interval integrations/s); largest relative defect of the integrated nodes {worst:.1e}; "
          f"exact Hessian blocks {tuple(hes.H.shape)} in {1e3 * t_hes:.2f} ms; second derivatives of 256 scenarios: {tuple(D2.shape)}")
    assert worst < 1e-12
    # next cycle: carry the pulses over, shift the solved horizon, start from where this one ended
    nmpc.update_stim()
    x, _, pulse_width = nmpc.advance(X, pulse_width)
    x, pulse_width = x.contiguous(), pulse_width.contiguous()
print(f"mean force at the end of cycle {args.cycles}: {float(x[1].mean()):.2f} N, mean A {float(x[2].mean()):.1f} (rest {float(params[0]):.0f})")
