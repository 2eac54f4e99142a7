"""Node-number comparison with multi start (reference: examples/sensitivity/node_shooting_comparison.py, which holds only the
line "TODO : Build a node number comparison with multi start").  Built here on the batched path: how the integrated force
depends on the discretisation -- shooting nodes per pulse interval x RK steps per node x scheme -- for a POPULATION of starting
points (muscles whose parameters and initial states scatter around the defaults), each combination one launch.

For every (scheme, nodes per pulse interval, steps per node) the script integrates the same 1 s / 33 Hz fatigue train for all
starts and reports the largest force deviation over nodes and starts from a fine RK4 solution (10 nodes x 40 steps per pulse
interval), next to the launch time: the table a user reads to pick OcpFes' n_shooting and OdeSolver settings.

    python examples/sensitivity/node_shooting_comparison.py [--starts 4096]
"""
import argparse
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))

from cocofest_b200 import ModelMaker, fes  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--starts", type=int, default=4096)
args = ap.parse_args()

dev = torch.device("cuda", 0)
final_time, hz, N = 1, 33, 10
stim = [k / hz for k in range(hz)]
model = ModelMaker.create_model("ding2003_with_fatigue", stim_time=stim, sum_stim_truncation=N)
B = args.starts
gen = torch.Generator().manual_seed(0)
params = (torch.tensor(model.param_vector())[:, None] * (0.8 + 0.4 * torch.rand((fes.FES_NP, B), generator=gen, dtype=torch.float64))).to(dev)
rest = torch.tensor(model.standard_rest_values()[:, 0])
x0 = (rest[:, None] * (0.9 + 0.2 * torch.rand((5, B), generator=gen, dtype=torch.float64)))
x0[0] = 0.05 * torch.rand(B, generator=gen, dtype=torch.float64)       # Cn, F: a muscle that is not at rest
x0[1] = 20.0 * torch.rand(B, generator=gen, dtype=torch.float64)
x0 = x0.to(dev)


def run(scheme, nodes_per_pulse, steps):
    """Force at the pulse instants [hz + 1, B] and the launch time in ms."""
    ns = hz * nodes_per_pulse
    win = torch.tensor(fes.stim_windows(stim, ns, final_time, N)[0], device=dev)
    fes.ivp(model.kind, True, params, x0, win, final_time / ns, n_steps=steps, n_shooting=ns, scheme=scheme, trace="nodes")   # warm-up
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = fes.ivp(model.kind, True, params, x0, win, final_time / ns, n_steps=steps, n_shooting=ns, scheme=scheme, trace="nodes")
    torch.cuda.synchronize()
    ms = 1e3 * (time.perf_counter() - t0)
    return res.trace[::nodes_per_pulse, 1, :], ms


fine, _ = run(4, 10, 40)
print(f"{B} starts, {hz} Hz for {final_time} s; reference: RK4, 10 nodes x 40 steps per pulse interval")
print(f"{'scheme':6s} {'nodes/pulse':>11s} {'steps/node':>10s} {'n_shooting':>10s} {'max |dF| (N)':>13s} {'rel':>9s} {'ms':>8s}")
rows = []
for scheme in (1, 2, 4):
    for npp in (1, 2, 5):
        for steps in (1, 2, 5, 10, 20):
            F, ms = run(scheme, npp, steps)
            err = float((F - fine).abs().max())
            rel = err / float(fine.abs().max())
            rows.append((scheme, npp, steps, err))
            print(f"RK{scheme:<4d} {npp:11d} {steps:10d} {hz * npp:10d} {err:13.3e} {rel:9.1e} {ms:8.2f}")

# the discretisation error is a function of the STEP h = 1 / (hz nodes steps) alone: nodes only restart the window table
by_h = {}
for scheme, npp, steps, err in rows:
    by_h.setdefault((scheme, npp * steps), []).append(err)
spread = max(max(v) / max(min(v), 1e-300) for v in by_h.values() if len(v) > 1 and max(v) > 1e-9)
print(f"same step through different (nodes, steps) splits: errors agree within a factor {spread:.3f}")
rk4 = {npp * steps: err for scheme, npp, steps, err in rows if scheme == 4}
order = np.log2(rk4[50] / rk4[100])
print(f"observed order of RK4 between 50 and 100 steps per pulse interval: {order:.2f} "
      f"(coarser steps are not asymptotic: tauc = {model.tauc} s against h = {1e3 / (hz * 5):.1f} ms at 5 steps)")
assert 3.5 < order < 4.5 and spread < 1.5
